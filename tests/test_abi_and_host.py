"""CPU-side checks: the C-ABI library loads and exports every symbol include/b2stereo.h declares,
the host mirror of the reference interface (parse_config_file, matcher construction), the
no-fallback rule, and the pair sharding used by bench.py (world_size-2 gloo)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built_lib():
    from stereo_match_b200 import build
    return build.build()


def test_abi_exports_every_declared_symbol(built_lib):
    hdr = open(os.path.join(ROOT, "include", "b2stereo.h")).read()
    declared = set(re.findall(r"B2S_API\s+[\w\s\*]+?\b(b2s_\w+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = ctypes.CDLL(built_lib)
    for name in sorted(declared):
        assert hasattr(lib, name), "libb2stereo.so does not export %s" % name
    from stereo_match_b200 import _lib
    assert declared == set(_lib.PROTOTYPES), "ctypes prototypes and header disagree: %s" % (declared ^ set(_lib.PROTOTYPES))
    lib.b2s_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.b2s_version()


def test_abi_only_exports_b2s_symbols(built_lib):
    out = subprocess.check_output(["nm", "-D", "--defined-only", built_lib], text=True)
    syms = [ln.split()[-1] for ln in out.splitlines() if " T " in ln]
    assert syms and all(s.startswith("b2s_") for s in syms), syms


def test_library_is_sm100a_only(built_lib):
    out = subprocess.run(["cuobjdump", "-lelf", built_lib], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_parse_config_file(tmp_path):
    import stereo_match_b200 as sm
    d = sm.parse_config_file(None)
    assert d["window_size"] == 3 and d["num_disparities"] == 160 and d["mode"] == "P" and d["sgbm_mode"] == "SGBM"
    ini = tmp_path / "settings.ini"
    ini.write_text("[disparity]\nwindow_size = 5\nmin_disparity=0\nnum_disparities=128\nblock_size=5\n"
                   "disp12_max_diff=1\nuniqueness_ratio=15\nspeckle_window_size=100\nspeckle_range=2\n"
                   "pre_filter_cap=63\nlmbda=80000\nsigma=1.2\nsgbm_mode=HH\n")
    d = sm.parse_config_file(str(ini))
    assert d["window_size"] == 5 and d["num_disparities"] == 128 and d["speckle_window_size"] == 100
    assert d["sigma"] == 1.2 and d["sgbm_mode"] == "HH" and d["mode"] == "P"
    assert sm.parse_config_file(str(tmp_path / "missing.ini"))["block_size"] == 5


def test_matcher_parameters_follow_reference(built_lib):
    """stereo_vision.py:148-163: P1 = 8*3*ws^2, P2 = 32*3*ws^2; getters/setters like cv2."""
    import stereo_match_b200 as sm
    s = sm.parse_config_file(None)
    s.update(window_size=5, num_disparities=128, sgbm_mode="HH")
    m = sm.create_left_matcher(s)
    assert (m.getP1(), m.getP2(), m.getNumDisparities(), m.getMode()) == (600, 2400, 128, sm.STEREO_SGBM_MODE_HH)
    g = m.geometry(1242)
    assert (g["minX1"], g["width1"], g["ndirs"], g["invalid"], g["ftzero"]) == (128, 1114, 8, -16, 63)
    m.setMinDisparity(-8); m.setPreFilterCap(0); m.setP1(0); m.setP2(0); m.setDisp12MaxDiff(0); m.setUniquenessRatio(-1)
    g = m.geometry(200)
    assert (g["minX1"], g["maxX1"], g["invalid"], g["ftzero"], g["P1"], g["P2"], g["disp12MaxDiff"], g["uniquenessRatio"]) \
        == (120, 192, -144, 15, 2, 5, 1, 10)
    r = sm.createRightMatcher(sm.create_left_matcher(s))
    assert (r.getMinDisparity(), r.getUniquenessRatio(), r.getDisp12MaxDiff(), r.getSpeckleWindowSize()) == (-127, 0, 1000000, 0)
    # method="BM" (stereo_vision.py:165-171): BM right matcher over the mirrored range; the WLS factory re-parametrises
    # the left matcher like cv2.ximgproc.createDisparityWLSFilter does (restated from memory: unpinned)
    lb = sm.create_left_matcher(s, method="BM")
    rb = sm.createRightMatcher(lb)
    assert (rb.getMinDisparity(), rb.getNumDisparities(), rb.getBlockSize(), rb.getTextureThreshold(), rb.getUniquenessRatio(),
            rb.getDisp12MaxDiff()) == (-127, 128, 5, 0, 0, 1000000)
    f = sm.createDisparityWLSFilter(lb)
    assert (lb.getTextureThreshold(), lb.getUniquenessRatio(), lb.getDisp12MaxDiff(), lb.getSpeckleWindowSize()) == (0, 0, 1000000, 0)
    assert f.getDepthDiscontinuityRadius() == 2 and f.getLRCthresh() == 24 and f.getLambda() == 8000.0
    ls = sm.create_left_matcher(s)
    f = sm.createDisparityWLSFilter(ls)
    f.setLambda(80000); f.setSigmaColor(1.2)
    assert (ls.getUniquenessRatio(), ls.getDisp12MaxDiff(), ls.getSpeckleWindowSize()) == (0, 1000000, 0)
    assert (f.getDepthDiscontinuityRadius(), f.getLambda(), f.getSigmaColor()) == (3, 80000.0, 1.2)
    with pytest.raises(RuntimeError, match="Method not supported"):
        sm.create_left_matcher(s, method="ELAS")
    # the cv2 geometry rule agrees with the oracle's
    from oracle import port
    port.build()
    kw = dict(minDisparity=-8, numDisparities=128, blockSize=5, P1=0, P2=0, disp12MaxDiff=0, preFilterCap=0, uniquenessRatio=-1)
    og = port.sgbm_geometry(200, **kw)
    assert (og["minX1"], og["maxX1"], og["INVALID"], og["ftzero"], og["P1"], og["P2"]) == (120, 192, -144, 15, 2, 5)


def test_no_cpu_fallback(built_lib):
    """Without a CUDA device the product must fail loudly, never route to a CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    import stereo_match_b200 as sm
    l = np.zeros((32, 64), np.uint8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sm.StereoSGBM_create(numDisparities=16).compute(l, l)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sm.reprojectImageTo3D(l, np.eye(4))


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "stereo_match_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src, f
    # cv2 may only appear in the synthetic-input generator
    for f in ("engine.py", "ops.py", "stereo_vision.py", "_lib.py", "batch.py"):
        p = os.path.join(ROOT, "stereo_match_b200", f)
        if os.path.exists(p):
            assert "import cv2" not in open(p).read(), f


def test_shard_ranges():
    from stereo_match_b200.batch import shard_range
    for n in (0, 1, 7, 8192):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


_GLOO_WORKER = r"""
import os, sys, json
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
from stereo_match_b200.batch import shard_range, gather_host_results
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 11
lo, hi = shard_range(n, rank, world)
# stand-in for the per-rank engine output: a function of the global pair index only
local = torch.stack([torch.full((4, 6), i, dtype=torch.int16) for i in range(lo, hi)])
full = gather_host_results(local, n, dst=0)
t = torch.tensor([float(rank + 1)], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)          # bench.py's max-over-ranks timing reduction
if rank == 0:
    ok = full.shape == (n, 4, 6) and all(int(full[i, 0, 0]) == i for i in range(n)) and t.item() == world
    print(json.dumps({"ok": bool(ok)}))
dist.destroy_process_group()
"""


def test_two_rank_sharding_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29571", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert '{"ok": true}' in r.stdout, r.stdout + r.stderr[-2000:]


def test_write_ply_matches_reference_writer(tmp_path):
    """Bytes of io_functions.write_ply (reference, run here by tests/golden/make_golden.py) for the same vertices."""
    import stereo_match_b200 as sm
    g = np.load(os.path.join(ROOT, "tests", "golden", "cv2_4_13_0.npz"))
    fn = tmp_path / "cloud.ply"
    sm.write_ply(str(fn), g["ply_verts"], g["ply_colors"])
    assert fn.read_bytes() == g["ply_bytes"].tobytes()
