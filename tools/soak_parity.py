"""Randomised parity soak of the engine against live cv2: many more shapes / parameters than the test suite (numDisparities 128 /
256, every mode, ratios up to 99, penalties up to the edge of the 16-bit domain, fully negative ranges, colour, batches).
usage: python tools/soak_parity.py [n_cases] [seed0] [sgbm|wide|big|params|struct|large|bm|bmstruct|ops|chain|misc|errors]      (GPU box; prints failures, exits 1 on any)
sgbm: the numDisparities 128 / 256 kernels (soak_cases.gen); wide: every regime (gen_wide); big: numDisparities up to 1024,
blocks up to 15 (gen_big); params: cv2's silent fallbacks and odd values (gen_params); large: realistic sizes, small batches (gen_large); struct / bmstruct: tie-heavy structured images
(constant, edges, checkerboards, exact shifts); bm: StereoBM (gen_bm); ops: remap, cvtColor,
medianBlur, filterSpeckles, reprojectImageTo3D; chain: the fused per-pair chain (sm.pipeline) against the separate cv2 calls;
misc: compute_disparity, project_points_3D, point_cloud, map2Dto3D; errors: invalid calls (both sides must raise, or neither).

Round 2: 600 cases found 25 mismatches, all three of them cv2 rules the restatement did not have (tests/vectors.py
SGBM_SOAK_VECTORS); 0 of 600 after."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cv2
import numpy as np
import torch

from soak_cases import gen, gen_wide, gen_big, gen_params, gen_struct, gen_large, gen_bm, gen_bm_struct, gen_chain, ops_failures

import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair

n, seed0 = (int(sys.argv[1]) if len(sys.argv) > 1 else 300), (int(sys.argv[2]) if len(sys.argv) > 2 else 0)
which = sys.argv[3] if len(sys.argv) > 3 else "sgbm"
bad = ran = 0


def soak_misc(k):
    """The reference-level functions: compute_disparity (left / right maps), project_points_3D, the ordered point cloud,
    map2Dto3D -- against cv2 and numpy directly.  -> names of what differs"""
    rng = np.random.default_rng(838300 + k)
    fails = []
    # compute_disparity (stereo_vision.py:132-184): the maps of both matchers for random settings
    D = 16 * int(rng.integers(1, 9))
    ws = int(rng.choice([1, 3, 5]))
    method = "BM" if rng.random() < 0.3 else "SGBM"
    bs = int(rng.choice([5, 9, 15, 21])) if method == "BM" else int(rng.choice([1, 3, 5, 7]))
    st = dict(window_size=ws, min_disparity=0 if method == "BM" else int(rng.choice([0, 0, 4])), num_disparities=D, block_size=bs,
              disp12_max_diff=int(rng.choice([-1, 1, 5])), uniqueness_ratio=int(rng.choice([0, 10, 15])),
              speckle_window_size=int(rng.choice([0, 50])), speckle_range=int(rng.choice([1, 2])), pre_filter_cap=int(rng.choice([15, 63])),
              lmbda=80000.0, sigma=1.2, sgbm_mode=str(rng.choice(["SGBM", "HH", "HH4"])))
    w, h = D + st["min_disparity"] + int(rng.integers(bs + 2, 120)), int(rng.integers(bs + 2, 70))
    l, r = synth_pair(w, h, D, seed=k)
    displ, filtered, dispr = sm.compute_disparity(l, r, st, method, return_right=True)
    if method == "SGBM":
        kwl = dict(minDisparity=st["min_disparity"], numDisparities=D, blockSize=bs, P1=24 * ws * ws, P2=96 * ws * ws, preFilterCap=st["pre_filter_cap"],
                   mode={"SGBM": 0, "HH": 1, "HH4": 3}[st["sgbm_mode"]])
        # createDisparityWLSFilter(left_matcher) switches the left matcher's own filters off (stereo_vision.py:172)
        wl = cv2.StereoSGBM_create(disp12MaxDiff=1000000, uniquenessRatio=0, speckleWindowSize=0, speckleRange=st["speckle_range"], **kwl).compute(l, r)
        kwr = dict(kwl, minDisparity=-(st["min_disparity"] + D) + 1)
        wr = cv2.StereoSGBM_create(disp12MaxDiff=1000000, uniquenessRatio=0, speckleWindowSize=0, speckleRange=0, **kwr).compute(r, l)
    else:
        ml = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
        ml.setDisp12MaxDiff(1000000); ml.setSpeckleWindowSize(0); ml.setTextureThreshold(0); ml.setUniquenessRatio(0)
        wl = ml.compute(l, r)
        mr = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
        mr.setMinDisparity(-D + 1); mr.setTextureThreshold(0); mr.setUniquenessRatio(0); mr.setDisp12MaxDiff(1000000); mr.setSpeckleWindowSize(0)
        wr = mr.compute(r, l)
    if not np.array_equal(displ, wl):
        fails.append("compute_disparity displ %s" % method)
    if not np.array_equal(dispr, wr):
        fails.append("compute_disparity dispr %s" % method)
    if filtered.dtype != np.int16 or filtered.shape != l.shape:
        fails.append("compute_disparity filtered shape")
    # project_points_3D (stereo_vision.py:187-210)
    Q = np.array([[1, 0, 0, -w / 2], [0, -1, 0, h / 2], [0, 0, 0, -float(rng.uniform(200, 1200))], [0, 0, float(rng.uniform(0.5, 5)), 0]])
    want = cv2.reprojectImageTo3D(displ.astype(np.float32) / 16.0, Q.astype(np.float32))
    got = sm.project_points_3D(displ, Q)
    nan = np.isnan(got) & np.isnan(want)
    if not np.array_equal(np.where(nan, 0, got.view(np.uint32)), np.where(nan, 0, want.view(np.uint32))):
        fails.append("project_points_3D")
    # the ordered point cloud (disparity_calculation.py:309-321)
    lo, hi = float(rng.choice([-np.inf, 0, 16 * 3])), float(rng.choice([np.inf, 16 * D // 2]))
    bgr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    for dd in (displ, displ.astype(np.float32) / 16):
        lo_, hi_ = (lo, hi) if dd.dtype == np.int16 else (lo / 16, hi / 16)
        mask = np.logical_and(dd < hi_, dd > lo_)
        pts = want[mask].copy()
        pts[np.isnan(pts)] = 0
        pts[np.isinf(pts)] = 0
        v, c = sm.point_cloud(dd, want, bgr, lo_, hi_)
        if not (np.array_equal(v.view(np.uint32), pts.view(np.uint32)) and np.array_equal(c, bgr[..., ::-1][mask])):
            fails.append("point_cloud %s" % dd.dtype)
    # map2Dto3D (mapTo3D.py:35-56)
    dt = rng.choice([np.uint8, np.uint16, np.int16, np.int32, np.float32, np.float64])
    gray = rng.integers(0, 256, (h, w), dtype=np.uint8)
    depth = (rng.random((h, w)) * float(rng.choice([1.0, 200.0, 30000.0]))).astype(dt)
    cx, cy, fx, fy = float(rng.uniform(0, w)), float(rng.uniform(0, h)), float(rng.uniform(100, 900)), float(rng.uniform(100, 900))
    rows, cols = np.where(gray > -1)
    z = depth[gray > -1]
    wantm = np.stack([-np.reshape((cols - cx) * z / fx, gray.shape), np.reshape((rows - cy) * z / fy, gray.shape), np.reshape(z, gray.shape)], axis=2)
    gotm = sm.map2Dto3D(gray, depth, cx, cy, fx, fy, -1)
    if not np.array_equal(gotm.view(np.uint64), wantm.astype(np.float64).view(np.uint64)):
        fails.append("map2Dto3D %s" % np.dtype(dt).name)
    return fails


def soak_errors(k):
    """Invalid calls: cv2 raises cv2.error -- the engine must raise too (and must not raise where cv2 accepts).
    -> descriptions of the disagreements"""
    rng = np.random.default_rng(949400 + k)
    out = []

    def both(name, f_cv2, f_sm):
        try:
            f_cv2()
            a = None
        except (cv2.error, TypeError, ValueError, SystemError, OverflowError) as ex:
            a = type(ex).__name__
        try:
            f_sm()
            b = None
        except (sm.error, TypeError, ValueError, OverflowError) as ex:
            b = type(ex).__name__
        if (a is None) != (b is None):
            out.append("%s: cv2 %s, engine %s" % (name, a or "accepts", b or "accepts"))

    h, w = int(rng.integers(20, 60)), int(rng.integers(80, 160))
    l = rng.integers(0, 256, (h, w), dtype=np.uint8)
    r = rng.integers(0, 256, (h, w), dtype=np.uint8)
    D = 16 * int(rng.integers(1, 4))
    bad_pairs = {
        "height mismatch": (l, r[:-1]), "width mismatch": (l, np.ascontiguousarray(r[:, :-1])), "float32 images": (l.astype(np.float32), r.astype(np.float32)),
        "uint16 images": (l.astype(np.uint16), r.astype(np.uint16)), "int8 images": (l.astype(np.int8), r.astype(np.int8)),
        # (2- and 4-channel pairs are left out: cv2 4.13.0 accepts them and then corrupts its heap in MODE_HH4; the engine takes 1 or 3)
        "gray vs colour": (l, np.stack([r] * 3, -1)), "empty": (l[:0], r[:0]), "one column": (l[:, :1], r[:, :1]),
        "too narrow": (np.ascontiguousarray(l[:, :D]), np.ascontiguousarray(r[:, :D])), "colour pair": (np.stack([l] * 3, -1), np.stack([r] * 3, -1)),
        "non-contiguous": (l[:, ::2], r[:, ::2]), "transposed view": (l.T, r.T), "different dtypes": (l, r.astype(np.float32)),
    }
    name = list(bad_pairs)[int(rng.integers(0, len(bad_pairs)))]
    a, b = bad_pairs[name]
    # (MODE_SGBM_3WAY is left out: cv2 only refuses W - maxD <= 0 there, the engine keeps the other modes' test)
    kw = dict(minDisparity=0, numDisparities=D, blockSize=5, P1=100, P2=400, mode=int(rng.choice([0, 1, 3])))
    both("SGBM " + name, lambda: cv2.StereoSGBM_create(**kw).compute(a, b), lambda: sm.StereoSGBM_create(**kw).compute(a, b))
    both("BM " + name, lambda: cv2.StereoBM_create(numDisparities=D, blockSize=9).compute(a, b),
         lambda: sm.StereoBM_create(numDisparities=D, blockSize=9).compute(a, b))
    # matcher parameters
    nd = int(rng.choice([0, -16, 8, 17, 24, 16, 48]))
    bsz = int(rng.choice([0, 3, 4, 5, 6, 21, 255, 257, -5, h | 1, (h + 2) | 1, 2 * h + 1]))
    both("BM numDisparities %d blockSize %d" % (nd, bsz), lambda: cv2.StereoBM_create(numDisparities=nd, blockSize=bsz).compute(l, r),
         lambda: sm.StereoBM_create(numDisparities=nd, blockSize=bsz).compute(l, r))
    nds = int(rng.choice([0, -16, 1, 7, 17, 16]))
    both("SGBM numDisparities %d" % nds, lambda: cv2.StereoSGBM_create(minDisparity=0, numDisparities=nds, blockSize=5).compute(l, r),
         lambda: sm.StereoSGBM_create(minDisparity=0, numDisparities=nds, blockSize=5).compute(l, r))
    for setter, val in (("PreFilterCap", int(rng.choice([0, 1, 31, 63, 64, 100]))), ("TextureThreshold", int(rng.choice([-1, 0, 10]))),
                        ("UniquenessRatio", int(rng.choice([-1, 0, 15, 101]))), ("PreFilterSize", int(rng.choice([3, 5, 9, 256, 4])))):
        def mk(mod, setter=setter, val=val):
            m = mod.StereoBM_create(numDisparities=32, blockSize=9)
            getattr(m, "set" + setter)(val)
            return m.compute(l, r)
        if hasattr(sm.StereoBM_create(numDisparities=32, blockSize=9), "set" + setter):
            both("BM set%s(%d)" % (setter, val), lambda: mk(cv2), lambda: mk(sm))
    # the stage operators
    d = rng.integers(-16, 500, (h, w)).astype(np.int16)
    Q = np.eye(4)
    badq = {"Q 3x4": Q[:3], "Q 4x3": Q[:, :3], "Q int": Q.astype(np.int32), "Q float32": Q.astype(np.float32), "Q 1-d": Q.ravel()}
    qn = list(badq)[int(rng.integers(0, len(badq)))]
    both("reproject " + qn, lambda: cv2.reprojectImageTo3D(d, badq[qn]), lambda: sm.reprojectImageTo3D(d, badq[qn]))
    # (a 3-d disparity array is a batch for the engine; cv2 returns None for an empty one)
    badd = {"float64 disparity": d.astype(np.float64), "uint16 disparity": d.astype(np.uint16), "int8 disparity": d.astype(np.int8)}
    dn = list(badd)[int(rng.integers(0, len(badd)))]
    both("reproject " + dn, lambda: cv2.reprojectImageTo3D(badd[dn], Q), lambda: sm.reprojectImageTo3D(badd[dn], Q))
    mx = rng.random((h, w)).astype(np.float32) * w
    my = rng.random((h, w)).astype(np.float32) * h
    badm = {"float64 maps": (mx.astype(np.float64), my.astype(np.float64)), "map size mismatch": (mx, my[:-1]), "int maps": (mx.astype(np.int32), my.astype(np.int32)),
            "smaller maps": (mx[:5, :7].copy(), my[:5, :7].copy())}
    mn = list(badm)[int(rng.integers(0, len(badm)))]
    both("remap " + mn, lambda: cv2.remap(l, badm[mn][0], badm[mn][1], cv2.INTER_LINEAR), lambda: sm.remap(l, badm[mn][0], badm[mn][1], sm.INTER_LINEAR))
    both("medianBlur float", lambda: cv2.medianBlur(d.astype(np.float64), 3), lambda: sm.medianBlur(d.astype(np.float64), 3))
    both("filterSpeckles uint8", lambda: cv2.filterSpeckles(l.copy(), 0, 10, 1), lambda: sm.filterSpeckles(l.copy(), 0, 10, 1))
    return out


for k in range(seed0, seed0 + n):
    if which == "errors":
        f = soak_errors(k)
        ran += 1
        if f:
            bad += 1
            print("FAIL errors case %d: %s" % (k, "; ".join(f)))
        continue
    if which == "misc":
        f = soak_misc(k)
        ran += 1
        if f:
            bad += 1
            print("FAIL misc case %d: %s" % (k, ", ".join(f)))
        continue
    if which == "ops":
        f = ops_failures(k, sm, cv2)
        ran += 1
        if f:
            bad += 1
            print("FAIL ops case %d: %s" % (k, ", ".join(f)))
        continue
    if which == "chain":
        c = gen_chain(k)
        if c is None:
            continue
        try:
            ref = cv2.StereoSGBM_create(**c["kw"])
            rl = cv2.remap(c["left"], c["maps_left"][0], c["maps_left"][1], cv2.INTER_LINEAR)
            rr = cv2.remap(c["right"], c["maps_right"][0], c["maps_right"][1], cv2.INTER_LINEAR)
            if c["cn"] == 3:
                rl, rr = cv2.cvtColor(rl, cv2.COLOR_BGR2GRAY), cv2.cvtColor(rr, cv2.COLOR_BGR2GRAY)
            wd = ref.compute(rl, rr)
        except cv2.error:
            continue
        wx = cv2.reprojectImageTo3D(wd if c["scale"] == 1.0 else wd.astype(np.float32) * np.float32(c["scale"]), c["Q"],
                                    handleMissingValues=c["hm"])
        m = sm.StereoSGBM_create(**c["kw"])
        disp, xyz, grl, grr = sm.pipeline(m, c["left"], c["right"], c["maps_left"], c["maps_right"], c["Q"], disp_scale=c["scale"],
                                          handleMissingValues=c["hm"])
        if m.overflowed():
            continue
        ran += 1
        f = []
        if not (np.array_equal(grl, rl) and np.array_equal(grr, rr)):
            f.append("rectified")
        if not np.array_equal(disp, wd):
            f.append("disparity")
        nan = np.isnan(xyz) & np.isnan(wx)
        if not np.array_equal(np.where(nan, 0, xyz.view(np.uint32)), np.where(nan, 0, wx.view(np.uint32))):
            f.append("xyz")
        if f:
            bad += 1
            print("FAIL chain case %d: %s, %r cn %d scale %g hm %d shape %r" % (k, ", ".join(f), c["kw"], c["cn"], c["scale"], c["hm"], c["left"].shape))
        continue
    if which in ("bm", "bmstruct"):
        D, bs, setters, l, r = (gen_bm if which == "bm" else gen_bm_struct)(k)
        ref = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
        m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
        for name, val in setters.items():
            getattr(ref, "set" + name[0].upper() + name[1:])(val)
            getattr(m, "set" + name[0].upper() + name[1:])(val)
        want = ref.compute(l, r)
        if "speckleWindowSize" in setters and not (want != (setters["minDisparity"] - 1) * 16).any():
            continue              # cv2 returns uninitialised memory when nothing is valid and a speckle window is set
        got = m.compute(l, r)
        ran += 1
        if not np.array_equal(got, want):
            bad += 1
            print("FAIL bm case %d: %d pixels differ, D %d bs %d %r shape %r" % (k, int((got != want).sum()), D, bs, setters, l.shape))
        continue
    c = {"sgbm": gen, "wide": gen_wide, "big": gen_big, "params": gen_params, "struct": gen_struct, "large": gen_large}[which](k)
    if c is None:
        continue
    kw, l, r = c[:3]
    cn, B = (c[3], c[4]) if which in ("sgbm", "large") else (l.ndim == 3 and 3 or 1, 1 + 2 * (k % 3 == 0))
    try:
        want = cv2.StereoSGBM_create(**kw).compute(l, r)
    except cv2.error:
        try:
            sm.StereoSGBM_create(**kw).compute(l, r)
        except sm.error:
            continue
        bad += 1
        print("FAIL case %d: cv2 raises, the engine does not, %r shape %r" % (k, kw, l.shape))
        continue
    try:
        m = sm.StereoSGBM_create(**kw)
        m.compute(l, r)
    except sm.error as ex:
        bad += 1
        print("FAIL case %d: the engine raises (%s), cv2 does not, %r shape %r" % (k, str(ex)[:80], kw, l.shape))
        continue
    if B == 1:
        got = m.compute(l, r)
    else:
        L = torch.from_numpy(np.stack([l] * B)).cuda()
        R = torch.from_numpy(np.stack([r] * B)).cuda()
        out = m.compute(L, R).cpu().numpy()
        got = out[B - 1]
        if not all(np.array_equal(out[i], out[0]) for i in range(B)):
            print("BATCH MISMATCH", k, kw, l.shape)
            bad += 1
    if m.overflowed():
        continue
    ran += 1
    if not np.array_equal(got, want):
        bad += 1
        print("FAIL case %d: %d pixels differ, %r shape %r cn %d batch %d" % (k, int((got != want).sum()), kw, l.shape, cn, B))
print("soak %s: %d cases run, %d failures" % (which, ran, bad))
sys.exit(1 if bad else 0)
