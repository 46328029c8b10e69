"""Pins the C oracle (oracle/stereo_oracle.c): bit-exact against live cv2 4.13.0 on every parity
vector of SURVEY.md section 4, and against the committed golden fixtures (tests/golden/).  CPU only."""
import os

import numpy as np
import pytest

from vectors import (SGBM_VECTORS, SGBM_ERROR_VECTORS, SGBM_COLOR_VECTORS, SGBM_SOAK_VECTORS, BM_VECTORS, sgbm_kwargs,
                     sgbm_pair, color_pair, soak_pair)
from stereo_match_b200.synth import synth_pair, KITTI_Q

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv2_4_13_0.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _ids(vs):
    return ["-".join(str(x) for x in v[:5]) + ("-m%d" % v[10] if len(v) > 10 else "") for v in vs]


@pytest.mark.parametrize("i,v", list(enumerate(SGBM_VECTORS)), ids=_ids(SGBM_VECTORS))
def test_sgbm_port_equals_cv2(i, v, port, cv2ref):
    l, r = sgbm_pair(i, v)
    kw = sgbm_kwargs(v)
    got, want = port.sgbm_compute(l, r, **kw), cv2ref.sgbm_compute(l, r, **kw)
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


@pytest.mark.parametrize("v", SGBM_SOAK_VECTORS, ids=["%s%d-" % v[:2] + "-".join(str(x) for x in v[2:7]) + "-m%d" % v[12] for v in SGBM_SOAK_VECTORS])
def test_sgbm_port_soak_regressions(v, port, cv2ref):
    """cv2 rules found by the randomised soaks: 3WAY's 8-lane winner on tied minima, its 16-bit wrapping uniqueness
    threshold, and "every S saturated -> no winner".  Each vector failed before the rule was restated."""
    l, r = soak_pair(v)
    kw = sgbm_kwargs(v[2:])
    got, want = port.sgbm_compute(l, r, **kw), cv2ref.sgbm_compute(l, r, **kw)
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


@pytest.mark.parametrize("i,seed", SGBM_COLOR_VECTORS)
def test_sgbm_port_equals_cv2_color(i, seed, port):
    """3-channel input: every colour channel contributes its own Sobel and raw BT cost."""
    import cv2
    v = SGBM_VECTORS[i]
    kw = sgbm_kwargs(v)
    l, r = color_pair(v[0], v[1], max(v[2] + v[3], 8), seed)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    got = port.sgbm_compute(l, r, **kw)
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


@pytest.mark.parametrize("v", SGBM_ERROR_VECTORS, ids=_ids(SGBM_ERROR_VECTORS))
def test_sgbm_port_rejects_like_cv2(v, port, cv2ref):
    import cv2
    l, r = sgbm_pair(0, v)
    with pytest.raises(cv2.error):
        cv2ref.sgbm_compute(l, r, **sgbm_kwargs(v))
    with pytest.raises(ValueError):
        port.sgbm_compute(l, r, **sgbm_kwargs(v))


def test_sgbm_port_cfg1_fullsize(port, cv2ref):
    from vectors import CFG1
    l, r = synth_pair(CFG1["w"], CFG1["h"], 64, seed=CFG1["seed"])
    assert np.array_equal(port.sgbm_compute(l, r, **CFG1["kw"]), cv2ref.sgbm_compute(l, r, **CFG1["kw"]))


def test_sgbm_paths_sum_to_S(port):
    """Direction decomposition: S = min(32767, sum of the per-direction L_r)."""
    v = SGBM_VECTORS[3]
    l, r = sgbm_pair(3, v)
    kw = sgbm_kwargs(v)
    _, Cv, Sv, _ = port.sgbm_compute(l, r, want_volumes=True, **kw)
    g = port.sgbm_geometry(l.shape[1], **kw)
    tot = np.zeros(Cv.shape, np.int64)
    for k in range(8):
        tot += port.sgbm_path(Cv, g["P1"], g["P2"], k)
    assert np.array_equal(np.minimum(tot, 32767).astype(np.int16), Sv)


@pytest.mark.parametrize("i,v", list(enumerate(BM_VECTORS)), ids=_ids(BM_VECTORS))
def test_bm_port_equals_cv2(i, v, port, cv2ref):
    w, h, D, bs, setters = v
    l, r = synth_pair(w, h, D, seed=200 + i)
    assert np.array_equal(port.bm_compute(l, r, D, bs, **setters), cv2ref.bm_compute(l, r, D, bs, **setters))


def test_stage_ports_equal_cv2(port, cv2ref):
    rng = np.random.default_rng(7)
    img = rng.integers(-16, 2048, (37, 53)).astype(np.int16)
    assert np.array_equal(port.median3(img), cv2ref.median3(img))
    sp = (rng.integers(0, 40, (60, 80)) * 16).astype(np.int16)
    sp[rng.random(sp.shape) < 0.3] = -16
    for size, diff in ((0, 0), (10, 16), (100, 32), (400, 48)):
        assert np.array_equal(port.filter_speckles(sp, -16, size, diff), cv2ref.filter_speckles(sp, -16, size, diff))
    d16 = rng.integers(-16, 2048, (50, 70)).astype(np.int16)
    d16[0, :4] = 0
    for d in (d16, d16.astype(np.float32) / 16):
        for hm in (False, True):
            a, b = port.reproject(d, KITTI_Q, hm), cv2ref.reproject(d, KITTI_Q, hm)
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    src = rng.integers(0, 256, (60, 90, 3), dtype=np.uint8)
    mx = (rng.random((50, 80), dtype=np.float32) * 96 - 3).astype(np.float32)
    my = (rng.random((50, 80), dtype=np.float32) * 66 - 3).astype(np.float32)
    assert np.array_equal(port.remap(src, mx, my), cv2ref.remap(src, mx, my))
    assert np.array_equal(port.remap(src[..., 0].copy(), mx, my), cv2ref.remap(src[..., 0].copy(), mx, my))
    assert np.array_equal(port.bgr2gray(src), cv2ref.bgr2gray(src))


def test_reproject_dense_q_signed_zeros(port, cv2ref):
    """Bitwise, the sign of zeros and infinities included (NaN payloads excepted), for Q matrices beyond the reference's."""
    from vectors import reproject_q_cases
    for Q, d in reproject_q_cases():
        for Qm in (Q, Q.astype(np.float32)):
            for dd in (d, d.astype(np.float32) / 16, (d & 0xff).astype(np.uint8)):
                for hm in (False, True):
                    want, got = cv2ref.reproject(dd, Qm, hm), port.reproject(dd, Qm, hm)
                    nan = np.isnan(want) & np.isnan(got)
                    assert np.array_equal(np.where(nan, 0, got.view(np.uint32)), np.where(nan, 0, want.view(np.uint32)))


# ---- committed golden fixtures (no cv2 needed) ----------------------------------------------
def test_golden_sgbm(port, gold):
    ids = sorted(int(k[4:-2]) for k in gold.files if k.startswith("sgbm") and k.endswith("_d"))
    assert len(ids) >= 8
    for i in ids:
        v = tuple(int(x) for x in gold["sgbm%d_v" % i])
        assert v == SGBM_VECTORS[i]
        got = port.sgbm_compute(gold["sgbm%d_l" % i], gold["sgbm%d_r" % i], **sgbm_kwargs(v))
        assert np.array_equal(got, gold["sgbm%d_d" % i]), "golden sgbm vector %d" % i


def test_golden_sgbm_soak(port, gold):
    ids = sorted(int(k[4:-2]) for k in gold.files if k.startswith("soak") and k.endswith("_d"))
    assert len(ids) == len(SGBM_SOAK_VECTORS)
    for i in ids:
        v = SGBM_SOAK_VECTORS[i]
        assert tuple(int(x) for x in gold["soak%d_v" % i]) == v[1:]
        got = port.sgbm_compute(gold["soak%d_l" % i], gold["soak%d_r" % i], **sgbm_kwargs(v[2:]))
        assert np.array_equal(got, gold["soak%d_d" % i]), "golden soak vector %d" % i


def test_golden_sgbm_color(port, gold):
    ids = sorted(int(k[5:-2]) for k in gold.files if k.startswith("color") and k.endswith("_d"))
    assert len(ids) >= 4
    for i in ids:
        got = port.sgbm_compute(gold["color%d_l" % i], gold["color%d_r" % i], **sgbm_kwargs(SGBM_VECTORS[i]))
        assert np.array_equal(got, gold["color%d_d" % i]), "golden colour vector %d" % i


def test_golden_bm(port, gold):
    for i in (0, 1, 3):
        D, bs, cap, tex, uq = (int(x) for x in gold["bm%d_p" % i])
        got = port.bm_compute(gold["bm%d_l" % i], gold["bm%d_r" % i], D, bs, cap, tex, uq)
        assert np.array_equal(got, gold["bm%d_d" % i])


def test_golden_stages(port, gold):
    d16, Q = gold["rp_d"], gold["rp_Q"]
    assert np.array_equal(port.reproject(d16, Q).view(np.uint32), gold["rp_xyz"].view(np.uint32))
    assert np.array_equal(port.reproject(d16, Q, True).view(np.uint32), gold["rp_xyz_missing"].view(np.uint32))
    assert np.array_equal(port.reproject(d16.astype(np.float32) / 16.0, Q.astype(np.float32)).view(np.uint32),
                          gold["rp_xyz_f32"].view(np.uint32))
    assert np.array_equal(port.remap(gold["rm_src"], gold["rm_mx"], gold["rm_my"]), gold["rm_dst3"])
    assert np.array_equal(port.remap(np.ascontiguousarray(gold["rm_src"][..., 1]), gold["rm_mx"], gold["rm_my"]), gold["rm_dst1"])
    assert np.array_equal(port.bgr2gray(gold["rm_src"]), gold["gray"])
    for size, diff in ((0, 0), (10, 16), (100, 32)):
        assert np.array_equal(port.filter_speckles(gold["sp_in"], -16, size, diff), gold["sp_out_%d_%d" % (size, diff)])
    assert np.array_equal(port.median3(gold["med_in"]), gold["med_out"])


@pytest.mark.parametrize("seed", range(40))
def test_sgbm_port_random_sweep(seed, port):
    """Seeded random parameter / size / mode / channel combinations inside the guaranteed-parity domain."""
    import cv2
    from random_cases import random_case
    kw, l, r = random_case(seed)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    got = port.sgbm_compute(l, r, **kw)
    assert np.array_equal(got, want), "%d pixels differ for %r %r" % ((got != want).sum(), kw, l.shape)


@pytest.mark.parametrize("seed", range(30))
def test_bm_port_random_sweep(seed, port, cv2ref):
    from random_cases import random_bm_case
    D, bs, setters, l, r = random_bm_case(seed)
    assert np.array_equal(port.bm_compute(l, r, D, bs, **setters), cv2ref.bm_compute(l, r, D, bs, **setters)), (D, bs, setters, l.shape)


def test_golden_bm_ranges_and_speckle(port, gold):
    """StereoBM with the speckle filter and with minDisparity != 0 (the right-view matcher's range) against committed
    cv2 4.13.0 outputs: works without cv2."""
    keys = ("preFilterCap", "textureThreshold", "uniquenessRatio", "speckleWindowSize", "speckleRange", "minDisparity", "disp12MaxDiff")
    ids = sorted(int(k[3:-2]) for k in gold.files if k.startswith("bmx") and k.endswith("_d"))
    assert len(ids) >= 6
    for i in ids:
        p = [int(x) for x in gold["bmx%d_p" % i]]
        got = port.bm_compute(gold["bmx%d_l" % i], gold["bmx%d_r" % i], p[0], p[1], **dict(zip(keys, p[2:])))
        assert np.array_equal(got, gold["bmx%d_d" % i]), "golden BM vector %d" % i
