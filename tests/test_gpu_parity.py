"""GPU parity tests: the CUDA path (through the C ABI) against the oracle, bit-exact.

Checkers: oracle.port (C restatement, gives per-stage intermediates) and oracle.cv2ref
(live cv2 4.13.0).  Sizes are chosen so the CPU side finishes in seconds; the full
BASELINE.json shapes are in test_gpu_fullsize.py.
"""
import numpy as np
import pytest
import torch

import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair, KITTI_Q
from vectors import (SGBM_VECTORS, SGBM_ERROR_VECTORS, SGBM_UNPINNED_VECTORS, SGBM_COLOR_VECTORS, SGBM_SOAK_VECTORS, BM_VECTORS,
                     sgbm_kwargs, sgbm_pair, color_pair, soak_pair)

pytestmark = pytest.mark.gpu


def _ids(vs):
    return ["-".join(str(x) for x in v[:5]) + ("-m%d" % v[10] if len(v) > 10 else "") for v in vs]


@pytest.mark.parametrize("i,v", list(enumerate(SGBM_VECTORS)), ids=_ids(SGBM_VECTORS))
def test_sgbm_matches_oracle(i, v, port, cv2ref):
    l, r = sgbm_pair(i, v)
    kw = sgbm_kwargs(v)
    want = port.sgbm_compute(l, r, **kw)
    assert np.array_equal(want, cv2ref.sgbm_compute(l, r, **kw)), "oracle port disagrees with cv2"
    m = sm.StereoSGBM_create(**kw)
    got = m.compute(l, r)
    assert got.dtype == np.int16 and got.shape == l.shape
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


@pytest.mark.parametrize("v", SGBM_SOAK_VECTORS, ids=["%s%d-" % v[:2] + "-".join(str(x) for x in v[2:7]) + "-m%d" % v[12] for v in SGBM_SOAK_VECTORS])
def test_sgbm_soak_regressions(v, port, cv2ref):
    """The cv2 rules the randomised soaks found (tests/vectors.py SGBM_SOAK_VECTORS): MODE_SGBM_3WAY's 8-lane winner on tied
    minima and its wrapping 16-bit uniqueness threshold; a pixel whose S are all saturated has no winner.  A batch of the
    pair and its row-reversed copy checks the warp-wide slow path against pixels of another image next to it."""
    l, r = soak_pair(v)
    kw = sgbm_kwargs(v[2:])
    want = cv2ref.sgbm_compute(l, r, **kw)
    assert np.array_equal(port.sgbm_compute(l, r, **kw), want), "oracle port disagrees with cv2"
    m = sm.StereoSGBM_create(**kw)
    got = m.compute(l, r)
    assert not m.overflowed()
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()
    l2, r2 = np.ascontiguousarray(l[::-1]), np.ascontiguousarray(r[::-1])
    out = m.compute(torch.from_numpy(np.stack([l, l2])).cuda(), torch.from_numpy(np.stack([r, r2])).cuda()).cpu().numpy()
    assert np.array_equal(out[0], want)
    assert np.array_equal(out[1], cv2ref.sgbm_compute(l2, r2, **kw))


@pytest.mark.parametrize("i,seed", SGBM_COLOR_VECTORS)
def test_sgbm_color_matches_oracle(i, seed, port):
    """3-channel input (reference: try_try.py:69-81), every mode, against the restatement and live cv2."""
    import cv2
    v = SGBM_VECTORS[i]
    kw = sgbm_kwargs(v)
    l, r = color_pair(v[0], v[1], max(v[2] + v[3], 8), seed)
    want = port.sgbm_compute(l, r, **kw)
    assert np.array_equal(want, cv2.StereoSGBM_create(**kw).compute(l, r)), "oracle port disagrees with cv2"
    got = sm.StereoSGBM_create(**kw).compute(l, r)
    assert got.shape == l.shape[:2] and np.array_equal(got, want), "%d pixels differ" % (got != want).sum()
    if i in (0, 26):            # a colour batch equals its singles
        L = torch.from_numpy(np.stack([l, l[::-1].copy()])).cuda()
        R = torch.from_numpy(np.stack([r, r[::-1].copy()])).cuda()
        out = sm.StereoSGBM_create(**kw).compute(L, R)
        assert np.array_equal(out[0].cpu().numpy(), want)
        assert np.array_equal(out[1].cpu().numpy(), port.sgbm_compute(l[::-1].copy(), r[::-1].copy(), **kw))


@pytest.mark.parametrize("i", [1, 7, 11, 13, 15, 27, 29, 33])      # 27.. : MODE_HH4 (zeroed bottom cost rows)
def test_sgbm_volumes_match_oracle(i, port):
    """Stage taps: raw block cost and aggregated S against the restatement."""
    v = SGBM_VECTORS[i]
    l, r = sgbm_pair(i, v)
    kw = sgbm_kwargs(v)
    _, Cw, Sw, _ = port.sgbm_compute(l, r, want_volumes=True, **kw)
    m = sm.StereoSGBM_create(**kw)
    m.keep_volumes(True)
    m.compute(torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda())
    cost, S, g = m.debug_volumes(1, l.shape[0], l.shape[1])
    D = kw["numDisparities"]
    cost = cost[0, :, :, :D].cpu().numpy().astype(np.int32) & 0xffff
    S = S[0, :, :, :D].cpu().numpy()
    assert np.array_equal(cost + g["P2"], Cw.astype(np.int32)), "block cost differs"
    assert np.array_equal(S, Sw), "aggregated S differs"


def test_sgbm_batch_equals_singles(cv2ref):
    """Pairs are independent: a batched call equals per-pair calls (and cv2)."""
    kw = sgbm_kwargs(SGBM_VECTORS[3])
    pairs = [synth_pair(120, 70, 32, seed=s) for s in range(5)]
    L = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    R = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    m = sm.StereoSGBM_create(**kw)
    out = m.compute(L, R)
    assert isinstance(out, torch.Tensor) and out.is_cuda and out.shape == L.shape
    for k, (l, r) in enumerate(pairs):
        assert np.array_equal(out[k].cpu().numpy(), cv2ref.sgbm_compute(l, r, **kw))


def test_sgbm_idempotent_and_handle_reuse(cv2ref):
    """Same handle, different shapes and parameters in sequence; results never depend on history."""
    m = sm.StereoSGBM_create(**sgbm_kwargs(SGBM_VECTORS[0]))
    for i in (0, 7, 2, 0):
        v = SGBM_VECTORS[i]
        kw = sgbm_kwargs(v)
        for k, val in kw.items():
            getattr(m, "set" + k[0].upper() + k[1:])(val)
        l, r = sgbm_pair(i, v)
        assert np.array_equal(m.compute(l, r), cv2ref.sgbm_compute(l, r, **kw))


@pytest.mark.parametrize("v", SGBM_ERROR_VECTORS, ids=_ids(SGBM_ERROR_VECTORS))
def test_sgbm_rejects_what_cv2_rejects(v, cv2ref):
    l, r = sgbm_pair(0, v)
    import cv2
    with pytest.raises(cv2.error):
        cv2ref.sgbm_compute(l, r, **sgbm_kwargs(v))
    with pytest.raises(sm.error, match="too small"):
        sm.StereoSGBM_create(**sgbm_kwargs(v)).compute(l, r)


def test_sgbm_bad_inputs():
    m = sm.StereoSGBM_create(numDisparities=16, blockSize=5)
    l, r = synth_pair(64, 32, 16, seed=0)
    with pytest.raises(sm.error, match="left.size"):
        m.compute(l, r[:, :-1])
    with pytest.raises(sm.error, match="CV_8U"):
        m.compute(l.astype(np.int16), r.astype(np.int16))
    with pytest.raises(sm.error, match="unknown mode"):
        sm.StereoSGBM_create(numDisparities=16, mode=4).compute(l, r)
    with pytest.raises(sm.error):
        m.compute(np.dstack([l] * 4), np.dstack([r] * 4))     # 4 channels


@pytest.mark.parametrize("v", SGBM_UNPINNED_VECTORS, ids=_ids(SGBM_UNPINNED_VECTORS))
def test_sgbm_flags_unpinned_domain(v):
    l, r = sgbm_pair(9, v)
    m = sm.StereoSGBM_create(**sgbm_kwargs(v))
    m.compute(l, r)
    assert m.overflowed()
    m2 = sm.StereoSGBM_create(**sgbm_kwargs(SGBM_VECTORS[0]))
    m2.compute(*sgbm_pair(0, SGBM_VECTORS[0]))
    assert not m2.overflowed()


@pytest.mark.parametrize("i,v", list(enumerate(BM_VECTORS)), ids=_ids(BM_VECTORS))
def test_bm_matches_oracle(i, v, port, cv2ref):
    w, h, D, bs, setters = v
    l, r = synth_pair(w, h, D, seed=200 + i)
    want = cv2ref.bm_compute(l, r, D, bs, **setters)
    assert np.array_equal(want, port.bm_compute(l, r, D, bs, **setters))
    m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
    for k, val in setters.items():
        getattr(m, "set" + k[0].upper() + k[1:])(val)
    got = m.compute(l, r)
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


def test_bm_errors():
    l, r = synth_pair(64, 32, 16, seed=0)
    with pytest.raises(sm.error, match="divisible by 16"):
        sm.StereoBM_create(numDisparities=24, blockSize=5).compute(l, r)
    with pytest.raises(sm.error, match="odd"):
        sm.StereoBM_create(numDisparities=16, blockSize=6).compute(l, r)
    with pytest.raises(sm.error, match="CV_8UC1"):
        sm.StereoBM_create(numDisparities=16, blockSize=5).compute(np.dstack([l] * 3), np.dstack([r] * 3))


@pytest.mark.parametrize("shape", [(37, 53), (1, 1), (1, 9), (9, 1), (375, 1242)])
def test_median3(shape, cv2ref):
    rng = np.random.default_rng(1)
    img = rng.integers(-16, 2048, shape).astype(np.int16)
    assert np.array_equal(sm.medianBlur(img, 3), cv2ref.median3(img))


@pytest.mark.parametrize("size,diff", [(0, 0), (10, 16), (100, 32), (400, 48)])
def test_filter_speckles(size, diff, cv2ref, port):
    rng = np.random.default_rng(2)
    for s in range(4):
        img = (rng.integers(0, 40, (60, 80)) * 16).astype(np.int16)
        img[rng.random((60, 80)) < 0.3] = -16
        if s == 3:
            img[:] = 160            # one component covering the image
        want = cv2ref.filter_speckles(img, -16, size, diff)
        assert np.array_equal(want, port.filter_speckles(img, -16, size, diff))
        got = torch.from_numpy(img.copy()).cuda()
        sm.filterSpeckles(got, -16, size, diff)
        assert np.array_equal(got.cpu().numpy(), want)


def test_filter_speckles_batched(cv2ref):
    rng = np.random.default_rng(3)
    imgs = (rng.integers(0, 30, (3, 40, 50)) * 16).astype(np.int16)
    imgs[rng.random(imgs.shape) < 0.3] = -16
    got = torch.from_numpy(imgs.copy()).cuda()
    sm.filterSpeckles(got, -16, 20, 32)
    for k in range(3):   # components must not leak across images of a batch
        assert np.array_equal(got[k].cpu().numpy(), cv2ref.filter_speckles(imgs[k], -16, 20, 32))


def test_reproject_bitwise(cv2ref):
    rng = np.random.default_rng(4)
    d16 = rng.integers(-16, 2048, (375, 1242)).astype(np.int16)
    d16[0, :5] = 0                                           # W = 0 -> +-inf / NaN
    Qh = np.float32([[1, 0, 0, -360], [0, -1, 0, 640], [0, 0, 0, -1164], [0, 0, 1, 0]])   # disparity_calculation.py:293-299
    for Q in (KITTI_Q, Qh):
        for d in (d16, d16.astype(np.float32) / 16, d16.astype(np.int32), (d16 & 0xff).astype(np.uint8)):
            for hm in (False, True):
                want = cv2ref.reproject(d, Q, hm)
                got = sm.reprojectImageTo3D(d, Q, handleMissingValues=hm)
                assert got.dtype == np.float32 and got.shape == want.shape
                assert np.array_equal(got.view(np.uint32), want.view(np.uint32))   # bitwise, incl. inf/NaN
                # the stated tolerance of north_star: <= 1e-5 relative on finite entries
                fin = np.isfinite(want)
                assert np.all(np.abs(got[fin] - want[fin]) <= 1e-5 * np.maximum(np.abs(want[fin]), 1e-30))


def test_reproject_dense_q_signed_zeros(cv2ref):
    """Dense Q matrices with zeros of both signs (tests/vectors.py reproject_q_cases): bitwise, signs of 0 / inf included."""
    from vectors import reproject_q_cases
    for Q, d in reproject_q_cases():
        for Qm in (Q, Q.astype(np.float32)):
            for dd in (d, d.astype(np.float32) / 16, (d & 0xff).astype(np.uint8)):
                for hm in (False, True):
                    want, got = cv2ref.reproject(dd, Qm, hm), sm.reprojectImageTo3D(dd, Qm, handleMissingValues=hm)
                    nan = np.isnan(want) & np.isnan(got)
                    assert np.array_equal(np.where(nan, 0, got.view(np.uint32)), np.where(nan, 0, want.view(np.uint32)))


def test_project_points_3d_matches_reference_semantics(cv2ref):
    """stereo_vision.py:203-209: /16 then reprojectImageTo3D with a float32 Q."""
    rng = np.random.default_rng(5)
    d16 = rng.integers(-16, 2048, (60, 80)).astype(np.int16)
    want = cv2ref.reproject(d16.astype(np.float32) / 16.0, KITTI_Q.astype(np.float32))
    got = sm.project_points_3D(d16, KITTI_Q)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_remap_and_gray(cv2ref):
    rng = np.random.default_rng(6)
    mx, my = cv2ref.kitti_rect_maps()
    img3 = rng.integers(0, 256, (375, 1242, 3), dtype=np.uint8)
    img1 = np.ascontiguousarray(img3[..., 0])
    rx = (rng.random((200, 300), dtype=np.float32) * (1242 + 6) - 3).astype(np.float32)
    ry = (rng.random((200, 300), dtype=np.float32) * (375 + 6) - 3).astype(np.float32)
    for src in (img3, img1):
        assert np.array_equal(sm.remap(src, mx, my, sm.INTER_LINEAR), cv2ref.remap(src, mx, my))
        assert np.array_equal(sm.remap(src, rx, ry, sm.INTER_LINEAR), cv2ref.remap(src, rx, ry))
    assert np.array_equal(sm.cvtColor(img3, sm.COLOR_BGR2GRAY), cv2ref.bgr2gray(img3))
    batch = np.stack([img3, img3[::-1].copy()])
    out = sm.remap(batch, mx, my, sm.INTER_LINEAR)
    assert np.array_equal(out[1], cv2ref.remap(batch[1], mx, my))


def test_remap_gray_fused(cv2ref):
    """remap of a BGR image + BGR2GRAY in one kernel == the two cv2 calls (disparity_calculation.py:207-208, 286-287)."""
    rng = np.random.default_rng(16)
    mx, my = cv2ref.kitti_rect_maps()
    img3 = rng.integers(0, 256, (375, 1242, 3), dtype=np.uint8)
    want = cv2ref.bgr2gray(cv2ref.remap(img3, mx, my))
    assert np.array_equal(sm.remap_gray(img3, mx, my), want)
    batch = np.stack([img3, img3[:, ::-1].copy()])
    out = sm.remap_gray(batch, mx, my)
    assert np.array_equal(out[0], want)
    assert np.array_equal(out[1], cv2ref.bgr2gray(cv2ref.remap(batch[1], mx, my)))


@pytest.mark.parametrize("cn", [1, 3])
def test_pipeline_one_call(cv2ref, cn):
    """b2s_pipeline: remap x2 (+ gray) -> StereoSGBM -> reprojectImageTo3D against the same chain of cv2 calls."""
    from stereo_match_b200.synth import KITTI_Q
    rng = np.random.default_rng(17)
    mx, my = cv2ref.kitti_rect_maps()
    kw = dict(minDisparity=0, numDisparities=64, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, preFilterCap=63,
              uniquenessRatio=15, speckleWindowSize=100, speckleRange=2, mode=sm.STEREO_SGBM_MODE_HH)
    l, r = synth_pair(1242, 375, 64, seed=5)
    if cn == 3:
        l = np.stack([l, np.roll(l, 1, axis=1), l[::-1]], axis=-1).copy()
        r = np.stack([r, np.roll(r, 1, axis=1), r[::-1]], axis=-1).copy()
    m = sm.StereoSGBM_create(**kw)
    disp, xyz, rl, rr = sm.pipeline(m, l, r, (mx, my), (mx, my), KITTI_Q)
    wl, wr = cv2ref.remap(l, mx, my), cv2ref.remap(r, mx, my)
    if cn == 3:
        wl, wr = cv2ref.bgr2gray(wl), cv2ref.bgr2gray(wr)
    assert np.array_equal(rl, wl) and np.array_equal(rr, wr)
    wd = cv2ref.sgbm_compute(wl, wr, **kw)
    assert np.array_equal(disp, wd)
    wxyz = cv2ref.reproject(wd.astype(np.float32) / 16.0, KITTI_Q)
    assert np.array_equal(xyz.view(np.uint32), wxyz.view(np.uint32))
    disp2, none, _, _ = sm.pipeline(m, l, r, (mx, my), (mx, my), None)
    assert none is None and np.array_equal(disp2, wd)


def test_compute_disparity_surface(cv2ref, port):
    """The reference's compute_disparity (stereo_vision.py:132-184) with settings.ini values: (displ, filtered_img).

    cv2.ximgproc.createDisparityWLSFilter(left_matcher) re-parametrises the left matcher (uniquenessRatio 0,
    disp12MaxDiff 1000000, speckleWindowSize 0) BEFORE it computes; that rule, createRightMatcher's and the filter
    are UNPINNED restatements (no cv2.ximgproc here) -- what IS pinned: the two raw maps against plain cv2 with those
    parameters, and the filter kernels against oracle/wls_oracle.c."""
    settings = sm.parse_config_file(None)
    settings.update(window_size=5, num_disparities=64)
    l, r = synth_pair(200, 80, 64, seed=11)
    displ, filtered, dispr = sm.compute_disparity(l, r, settings, return_right=True)
    raw = dict(numDisparities=64, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1000000, uniquenessRatio=0,
               speckleWindowSize=0, preFilterCap=63)
    assert np.array_equal(displ, cv2ref.sgbm_compute(l, r, minDisparity=0, speckleRange=2, **raw))
    assert np.array_equal(dispr, cv2ref.sgbm_compute(r, l, minDisparity=-63, speckleRange=0, **raw))
    want, _ = port.wls_filter(displ, dispr, l, (64, 0, 200 - 64, 80), min_disp=0, lambda_=80000, sigma=1.2, radius=3)
    assert filtered.dtype == np.int16 and filtered.shape == l.shape and np.array_equal(filtered, want)
    two = sm.compute_disparity(l, r, settings)
    assert len(two) == 2 and np.array_equal(two[0], displ) and np.array_equal(two[1], filtered)
    with pytest.raises(RuntimeError, match="Method not supported"):
        sm.compute_disparity(l, r, settings, method="ELAS")
    # method="BM" (stereo_vision.py:165-166): StereoBM left matcher, BM right matcher over the mirrored range
    settings["num_disparities"] = 32
    displ_bm, filt_bm, dispr_bm = sm.compute_disparity(l, r, settings, method="BM", return_right=True)
    assert np.array_equal(displ_bm, cv2ref.bm_compute(l, r, 32, 5, textureThreshold=0, uniquenessRatio=0,
                                                       disp12MaxDiff=1000000, speckleWindowSize=0))
    assert np.array_equal(dispr_bm, cv2ref.bm_compute(r, l, 32, 5, minDisparity=-31, textureThreshold=0, uniquenessRatio=0,
                                                       disp12MaxDiff=1000000, speckleWindowSize=0))
    want_bm, _ = port.wls_filter(displ_bm, dispr_bm, l, (32 + 2, 2, 200 - 34 - 2, 80 - 4), min_disp=0, lambda_=80000,
                                 sigma=1.2, radius=2)
    assert np.array_equal(filt_bm, want_bm)


@pytest.mark.parametrize("w,h,D,bs,minD", [(120, 50, 32, 9, -31), (140, 40, 16, 7, -10), (140, 40, 16, 7, -40), (140, 40, 16, 7, 5),
                                            (140, 41, 48, 11, -47), (300, 80, 128, 15, -127), (90, 45, 64, 21, -63),
                                            (100, 40, 16, 5, -200), (160, 60, 32, 23, -31), (60, 40, 32, 5, -31)])
def test_bm_min_disparity_matches_cv2(w, h, D, bs, minD, cv2ref):
    """StereoBM with minDisparity != 0 (the right-view matcher of createRightMatcher uses -(D - 1)), incl. the columns
    where cv2 clamps the right window only (negative minDisparity) and the generic kernel (blockSize 23)."""
    l, r = synth_pair(w, h, max(D + minD, 8), seed=5)
    if minD < 0:
        l, r = r, l
    # (cv2 4.13.0 with minDisparity > 0 AND disp12MaxDiff >= 0 writes minDisparity stray pixels into the first row below
    #  the valid rectangle -- an overrun of its own, not a contract; the ximgproc factories only produce minD <= 0)
    for extra in ({}, dict(textureThreshold=0, uniquenessRatio=0, **(dict(disp12MaxDiff=1000000) if minD <= 0 else {}))):
        m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
        m.setMinDisparity(minD)
        for k, v in extra.items():
            getattr(m, "set" + k[0].upper() + k[1:])(v)
        want = cv2ref.bm_compute(l, r, D, bs, minDisparity=minD, **extra)
        got = m.compute(l, r)
        if minD > 0:        # cv2 also leaks a few computed values into the row below the valid rectangle
            want[h - bs // 2:] = (minD - 1) * 16
        assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()
    m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
    m.setDisp12MaxDiff(1)
    with pytest.raises(sm.error, match="disp12MaxDiff"):
        m.compute(l, r)


def test_launch_counter_and_library_loaded():
    from stereo_match_b200 import _lib
    before = _lib.load().b2s_launch_count()
    l, r = synth_pair(96, 64, 16, seed=0)
    sm.StereoSGBM_create(numDisparities=16, blockSize=5).compute(l, r)
    assert _lib.load().b2s_launch_count() - before >= 6
    with open("/proc/self/maps") as f:
        assert "libb2stereo.so" in f.read()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_filter_speckles_stress(seed, cv2ref):
    """Disparity-like maps at KITTI size (smooth surfaces, holes, noise), a serpentine component
    (long union-find chains) and a batch: exact agreement with cv2's flood fill."""
    rng = np.random.default_rng(50 + seed)
    H, W = 375, 1242
    yy, xx = np.mgrid[0:H, 0:W]
    base = (200 + 3 * yy + (xx // 97) * 40 + rng.integers(-8, 9, (H, W))).astype(np.int16)
    base[rng.random((H, W)) < 0.15] = -16
    base[rng.random((H, W)) < 0.02] = 3000
    snake = np.full((H, W), -16, np.int16)
    snake[::2, :] = 500
    snake[1::4, -1] = 500
    snake[3::4, 0] = 500
    imgs = np.stack([base, snake, (base // 4 * 4).astype(np.int16)])
    for size, diff in ((100, 32), (1000, 16), (400000, 48)):
        got = torch.from_numpy(imgs.copy()).cuda()
        sm.filterSpeckles(got, -16, size, diff)
        for k in range(len(imgs)):
            want = cv2ref.filter_speckles(imgs[k], -16, size, diff)
            assert np.array_equal(got[k].cpu().numpy(), want), (k, size, diff, int((got[k].cpu().numpy() != want).sum()))


def _cloud_numpy(disp, xyz, bgr, lo, hi):
    """disparity_calculation.py:309-321 restated with numpy's own boolean-mask indexing."""
    mask = np.logical_and(disp < hi, disp > lo)
    pts = xyz[mask].copy()
    pts[np.isnan(pts)] = 0
    pts[np.isinf(pts)] = 0
    return pts, (bgr[..., ::-1][mask] if bgr is not None else None)


@pytest.mark.parametrize("dtype", [np.int16, np.float32, np.uint8, np.int32])
def test_point_cloud_matches_numpy_mask(dtype):
    rng = np.random.default_rng(11)
    H, W = 53, 1031                                    # not a multiple of the block size, several blocks
    disp = rng.integers(-16, 40, (H, W)).astype(dtype) if dtype != np.uint8 else rng.integers(0, 40, (H, W)).astype(dtype)
    xyz = rng.standard_normal((H, W, 3)).astype(np.float32)
    xyz[rng.random((H, W)) < 0.05] = np.inf
    xyz[rng.random((H, W)) < 0.05, 1] = np.nan
    xyz[rng.random((H, W)) < 0.05, 2] = -np.inf
    bgr = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    for lo, hi in ((10, 20), (-np.inf, np.inf), (100, 200), (float(disp.min()), np.inf)):
        wp, wc = _cloud_numpy(disp, xyz, bgr, lo, hi)
        gp, gc = sm.point_cloud(disp, xyz, bgr, lo, hi)
        assert gp.shape == wp.shape and np.array_equal(gp.view(np.uint32), wp.view(np.uint32))
        assert np.array_equal(gc, wc)
    gp, gc = sm.point_cloud(disp, xyz, None, 10, 20)
    assert gc is None and np.array_equal(gp.view(np.uint32), _cloud_numpy(disp, xyz, None, 10, 20)[0].view(np.uint32))


def test_point_cloud_batch_and_chain(cv2ref):
    """Batch offsets; and the reference's tail (compute -> reprojectImageTo3D -> mask -> PLY vertices) on the GPU."""
    pairs = [synth_pair(160, 60, 32, seed=s) for s in range(3)]
    L = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    R = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    m = sm.StereoSGBM_create(minDisparity=0, numDisparities=32, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1,
                             uniquenessRatio=10, speckleWindowSize=50, speckleRange=2, preFilterCap=63)
    disp = m.compute(L, R)
    xyz = sm.reprojectImageTo3D(disp, KITTI_Q)
    bgr = torch.stack([L, R, L], -1).contiguous()
    verts, cols, offs = sm.point_cloud(disp, xyz, bgr, lo=100, hi=300)
    offs = offs.cpu().numpy()
    assert offs[0] == 0 and offs[-1] == verts.shape[0]
    for b in range(3):
        wp, wc = _cloud_numpy(disp[b].cpu().numpy(), xyz[b].cpu().numpy(), bgr[b].cpu().numpy(), 100, 300)
        assert np.array_equal(verts[offs[b]:offs[b + 1]].cpu().numpy().view(np.uint32), wp.view(np.uint32))
        assert np.array_equal(cols[offs[b]:offs[b + 1]].cpu().numpy(), wc)


def test_concurrent_matchers_on_two_streams():
    """One handle per (device, stream): two host threads, each with its own CUDA stream and its own matchers, compute
    interleaved cases at the same time; nothing in the library is shared between them (thread-local error / counters,
    per-handle workspaces), so every result equals cv2's."""
    import threading
    import cv2
    from random_cases import random_case, random_bm_case
    cases = [random_case(s) for s in (3, 12, 20, 33, 41, 57)]
    want = [cv2.StereoSGBM_create(**kw).compute(l, r) for kw, l, r in cases]
    D, bs, setters, bl, br = random_bm_case(4)
    ref = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
    for name, val in setters.items():
        getattr(ref, "set" + name[0].upper() + name[1:])(val)
    bm_want = ref.compute(bl, br)
    errs = []

    def worker(tid):
        try:
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                ms = [sm.StereoSGBM_create(**kw) for kw, _, _ in cases]
                bm = sm.StereoBM_create(numDisparities=D, blockSize=bs)
                for name, val in setters.items():
                    getattr(bm, "set" + name[0].upper() + name[1:])(val)
                dev = [(torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()) for _, l, r in cases]
                BL, BR = torch.from_numpy(bl).cuda(), torch.from_numpy(br).cuda()
                rng = np.random.default_rng(tid)
                for rep in range(25):
                    for j in rng.permutation(len(cases)):
                        got = ms[j].compute(*dev[j]).cpu().numpy()
                        if not np.array_equal(got, want[j]):
                            errs.append("thread %d rep %d case %d: %d pixels differ" % (tid, rep, j, (got != want[j]).sum()))
                    if not np.array_equal(bm.compute(BL, BR).cpu().numpy(), bm_want):
                        errs.append("thread %d rep %d: StereoBM differs" % (tid, rep))
        except Exception as ex:          # noqa: BLE001  (reported through the assertion below)
            errs.append("thread %d: %r" % (tid, ex))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errs, errs[:5]


def test_boundary_rules_found_by_the_errors_soak(cv2ref):
    """tools/soak_parity.py errors: cv2.StereoBM_create maps any numDisparities <= 0 to 64 and refuses a block that is not
    SMALLER than the image; cv2.filterSpeckles also takes CV_8UC1; MODE_SGBM / MODE_HH refuse a too narrow image, MODE_HH4 does not."""
    import cv2
    rng = np.random.default_rng(3)
    l = rng.integers(0, 256, (31, 100), dtype=np.uint8)
    r = rng.integers(0, 256, (31, 100), dtype=np.uint8)
    for nd in (-16, -1, 0):
        m = sm.StereoBM_create(numDisparities=nd, blockSize=9)
        assert m.getNumDisparities() == cv2.StereoBM_create(numDisparities=nd, blockSize=9).getNumDisparities() == 64
        assert np.array_equal(m.compute(l, r), cv2.StereoBM_create(numDisparities=nd, blockSize=9).compute(l, r))
    for bs, ok in ((29, True), (31, False), (33, False)):             # H = 31
        if ok:
            assert np.array_equal(sm.StereoBM_create(numDisparities=16, blockSize=bs).compute(l, r),
                                  cv2.StereoBM_create(numDisparities=16, blockSize=bs).compute(l, r))
        else:
            with pytest.raises(cv2.error):
                cv2.StereoBM_create(numDisparities=16, blockSize=bs).compute(l, r)
            with pytest.raises(sm.error, match="not larger than image width or height"):
                sm.StereoBM_create(numDisparities=16, blockSize=bs).compute(l, r)
    for t in range(20):
        img = (rng.integers(0, 6, (30, 40)) * int(rng.integers(1, 40))).astype(np.uint8)
        nv, size, diff = int(rng.choice([0, 255, 7, -1, 300])), int(rng.integers(0, 30)), int(rng.integers(0, 50))
        a, b = img.copy(), img.copy()
        cv2.filterSpeckles(a, nv, size, diff)
        out, _ = sm.filterSpeckles(b, nv, size, diff)
        assert out is b and b.dtype == np.uint8 and np.array_equal(a, b)
    narrow_l, narrow_r = np.ascontiguousarray(l[:, :34]), np.ascontiguousarray(r[:, :34])
    for mode in (0, 1):
        with pytest.raises(sm.error, match="too small"):
            sm.StereoSGBM_create(numDisparities=32, blockSize=5, P1=100, P2=400, mode=mode).compute(narrow_l, narrow_r)
    kw = dict(minDisparity=0, numDisparities=32, blockSize=5, P1=100, P2=400, mode=3)
    assert np.array_equal(sm.StereoSGBM_create(**kw).compute(narrow_l, narrow_r), cv2.StereoSGBM_create(**kw).compute(narrow_l, narrow_r))


@pytest.mark.parametrize("seed", range(80))
def test_sgbm_random_sweep(seed):
    """Seeded random parameter / size / mode / channel combinations against live cv2 (tests/random_cases.py)."""
    import cv2
    from random_cases import random_case
    kw, l, r = random_case(seed)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    m = sm.StereoSGBM_create(**kw)
    got = m.compute(l, r)
    assert not m.overflowed(), "case left the guaranteed domain: %r" % (kw,)
    assert np.array_equal(got, want), "%d pixels differ for %r %r" % ((got != want).sum(), kw, l.shape)


@pytest.mark.parametrize("seed", range(60))
def test_bm_random_sweep(seed, cv2ref):
    from random_cases import random_bm_case
    D, bs, setters, l, r = random_bm_case(seed)
    want = cv2ref.bm_compute(l, r, D, bs, **setters)
    m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
    for k, val in setters.items():
        getattr(m, "set" + k[0].upper() + k[1:])(val)
    got = m.compute(l, r)
    assert np.array_equal(got, want), "%d pixels differ for %r" % ((got != want).sum(), (D, bs, setters, l.shape))


@pytest.mark.parametrize("seed", range(12))
def test_stage_ops_random_sweep(seed, cv2ref):
    """remap (1 / 3 channels), cvtColor, medianBlur, filterSpeckles, reprojectImageTo3D on random shapes."""
    import cv2
    rng = np.random.default_rng(600 + seed)
    sh, sw = int(rng.integers(3, 90)), int(rng.integers(3, 150))
    H, W = int(rng.integers(1, 80)), int(rng.integers(1, 140))
    src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
    mx = (rng.random((H, W), dtype=np.float32) * (sw + 8) - 4).astype(np.float32)
    my = (rng.random((H, W), dtype=np.float32) * (sh + 8) - 4).astype(np.float32)
    assert np.array_equal(sm.remap(src, mx, my, sm.INTER_LINEAR), cv2.remap(src, mx, my, cv2.INTER_LINEAR))
    g = np.ascontiguousarray(src[..., 0])
    assert np.array_equal(sm.remap(g, mx, my, sm.INTER_LINEAR), cv2.remap(g, mx, my, cv2.INTER_LINEAR))
    assert np.array_equal(sm.cvtColor(src, sm.COLOR_BGR2GRAY), cv2.cvtColor(src, cv2.COLOR_BGR2GRAY))
    d = (rng.integers(-1, 60, (H, W)) * int(rng.integers(1, 17))).astype(np.int16)
    if min(H, W) >= 1:
        assert np.array_equal(sm.medianBlur(d, 3), cv2.medianBlur(d, 3))
    for size, diff in ((int(rng.integers(0, 60)), int(rng.integers(0, 40))),):
        a, b = d.copy(), d.copy()
        sm.filterSpeckles(a, -16, size, diff)
        cv2.filterSpeckles(b, -16, size, diff)
        assert np.array_equal(a, b), (size, diff, d.shape)
    Q = KITTI_Q.copy()
    Q[3, 3] = float(rng.random())
    for hm in (False, True):
        assert np.array_equal(sm.reprojectImageTo3D(d, Q, handleMissingValues=hm).view(np.uint32),
                              cv2.reprojectImageTo3D(d, Q, handleMissingValues=hm).view(np.uint32))


def _ref_map2Dto3D(pt2D, ptd, cx, cy, fx, fy, threshold):
    """The reference's own function body (mapTo3D.py:35-56), restated for the test."""
    rows, cols = np.where(pt2D > threshold)
    z = ptd[pt2D > threshold]
    u3d = (cols - cx) * z / fx
    v3d = (rows - cy) * z / fy
    return np.stack([-np.reshape(u3d, pt2D.shape), np.reshape(v3d, pt2D.shape), np.reshape(z, pt2D.shape)], axis=2)


@pytest.mark.parametrize("dt", [np.uint8, np.uint16, np.int16, np.int32, np.float32, np.float64])
def test_map2Dto3D_matches_reference_expression(dt):
    """mapTo3D.py:116: map2Dto3D(gray, depth, c_x, c_y, focal_x, focal_y, -1): bitwise float64."""
    rng = np.random.default_rng(11)
    gray = rng.integers(0, 256, (37, 53), dtype=np.uint8)
    depth = (rng.random((37, 53)) * 200).astype(dt)
    args = (319.5, 239.5, 525.0, 530.5, -1)
    want = _ref_map2Dto3D(gray, depth, *args)
    got = sm.map2Dto3D(gray, depth, *args)
    assert got.dtype == np.float64 and got.shape == (37, 53, 3)
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
    with pytest.raises(ValueError, match="cannot reshape array of size"):   # the reference fails the same way
        sm.map2Dto3D(gray, depth, 319.5, 239.5, 525.0, 530.5, 100)
    t = sm.map2Dto3D(torch.from_numpy(gray).cuda(), torch.from_numpy(depth).cuda() if dt != np.uint16 else
                     torch.from_numpy(depth.astype(np.int32)).cuda(), *args)
    assert t.is_cuda and np.array_equal(t.cpu().numpy().view(np.uint64), want.view(np.uint64))
