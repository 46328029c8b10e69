"""GPU tests of the disparity WLS filter (stereo_match_b200/csrc/wls.cu, stereo_match_b200/wls.py): the drop-in for
cv2.ximgproc.createDisparityWLSFilter(...).filter(displ, gray_l, None, dispr) of stereo_vision/stereo_vision.py:171-182.

PARITY UNPINNED against cv2 (no cv2.ximgproc on either box).  The bar here: the CUDA path, through the C ABI, is
BIT-EXACT with the restatement oracle/wls_oracle.c (tolerance 0 Q4 units; float32, same operation order, no fma
contraction), plus the size-independent properties of the algorithm at the KITTI shape."""
import numpy as np
import pytest
import torch

import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair

pytestmark = pytest.mark.gpu


def _case(seed, h, w, D, cn):
    rng = np.random.default_rng(seed)
    l, r = synth_pair(w, h, D, seed=seed)
    m = sm.StereoSGBM_create(minDisparity=0, numDisparities=D, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1000000,
                             preFilterCap=63, uniquenessRatio=0)
    mr = sm.createRightMatcher(m)
    dl, dr = m.compute(l, r), mr.compute(r, l)
    g = l if cn == 1 else np.stack([l, np.roll(l, 1, 0), rng.integers(0, 256, l.shape, dtype=np.uint8)], -1)
    return g, dl, dr


@pytest.mark.parametrize("seed,h,w,D,cn,conf", [(1, 40, 90, 16, 1, True), (2, 33, 130, 32, 3, True), (3, 70, 65, 16, 1, False),
                                                 (4, 1, 80, 16, 1, True), (5, 50, 49, 32, 3, False), (6, 97, 161, 64, 1, True)])
def test_wls_matches_restatement(seed, h, w, D, cn, conf, port):
    g, dl, dr = _case(seed, h, w, D, cn)
    f = sm.DisparityWLSFilter(conf, D if conf else 0, 0, 0, 0, 0)
    f.setLambda(8000.0 * seed)
    f.setSigmaColor(0.8 + 0.3 * seed)
    f.setDepthDiscontinuityRadius(1 + seed % 4)
    got = f.filter(dl, g, None, dr if conf else None)
    roi = (D, 0, w - D, h) if conf else (0, 0, w, h)
    want, wconf = port.wls_filter(dl, dr if conf else None, g, roi, use_confidence=conf, lambda_=8000.0 * seed,
                                  sigma=0.8 + 0.3 * seed, radius=1 + seed % 4)
    assert got.dtype == np.int16 and np.array_equal(got, want), "%d pixels differ (max %d Q4 units)" % (
        (got != want).sum(), np.abs(got.astype(int) - want).max())
    if conf:
        assert np.array_equal(f.getConfidenceMap().view(np.uint32), wconf.view(np.uint32))
        assert f.getROI() == roi


def test_wls_roi_argument_batch_and_torch(port):
    g, dl, dr = _case(9, 48, 120, 32, 1)
    f = sm.createDisparityWLSFilterGeneric(True)
    roi = (40, 5, 70, 30)
    got = f.filter(dl, g, None, dr, roi)
    want, _ = port.wls_filter(dl, dr, g, roi, lambda_=8000.0, sigma=1.0, radius=5)
    assert np.array_equal(got, want)
    # batch of device tensors == singles
    G = torch.from_numpy(np.stack([g, g[::-1].copy()])).cuda()
    DL = torch.from_numpy(np.stack([dl, dl[::-1].copy()])).cuda()
    DR = torch.from_numpy(np.stack([dr, dr[::-1].copy()])).cuda()
    out = f.filter(DL, G, None, DR, roi)
    assert out.is_cuda and np.array_equal(out[0].cpu().numpy(), want)
    want1, _ = port.wls_filter(dl[::-1].copy(), dr[::-1].copy(), g[::-1].copy(), roi, lambda_=8000.0, sigma=1.0, radius=5)
    assert np.array_equal(out[1].cpu().numpy(), want1)
    dst = np.zeros_like(dl)
    assert f.filter(dl, g, dst, dr, roi) is dst and np.array_equal(dst, want)


def test_wls_argument_errors():
    g, dl, dr = _case(10, 20, 60, 16, 1)
    f = sm.createDisparityWLSFilterGeneric(True)
    with pytest.raises(sm.error, match="disparity_map_right"):
        f.filter(dl, g)
    with pytest.raises(sm.error, match="CV_16S"):
        f.filter(dl.astype(np.float32), g, None, dr)
    with pytest.raises(sm.error, match="CV_8U"):
        f.filter(dl, g.astype(np.float32), None, dr)
    with pytest.raises(sm.error, match="rectangle"):
        f.filter(dl, g, None, dr, (30, 0, 60, 20))
    with pytest.raises(sm.error, match="StereoBM and StereoSGBM"):
        sm.createDisparityWLSFilter(object())


def test_wls_kitti_shape_properties():
    """cfg3 shape: fixed point of a constant map, value outside the rectangle, determinism, lambda -> 0."""
    h, w, D = 375, 1242, 128
    rng = np.random.default_rng(0)
    g = rng.integers(0, 256, (h, w), dtype=np.uint8)
    d = np.full((h, w), 40 * 16, np.int16)
    f = sm.DisparityWLSFilter(True, D, 0, 0, 0, 0)
    f.setLambda(80000)
    f.setSigmaColor(1.2)
    out = f.filter(d, g, None, (-d).astype(np.int16))
    assert np.all(out[:, D:] == 640) and np.all(out[:, :D] == -16)
    l, r = synth_pair(w, h, D, seed=1)
    settings = sm.parse_config_file(None)
    settings.update(window_size=5, num_disparities=D, sgbm_mode="HH")
    displ, filt, dispr = sm.compute_disparity(l, r, settings, return_right=True)
    again = sm.compute_disparity(l, r, settings)[1]
    assert np.array_equal(filt, again)
    valid = displ[:, D:] > 0
    assert np.median(np.abs(filt[:, D:].astype(int) - displ[:, D:])[valid]) <= 8     # stays close to the raw map where it is right
    f0 = sm.DisparityWLSFilter(False)
    f0.setLambda(0.0)
    assert np.array_equal(f0.filter(displ, l), displ)


def test_reference_script_lines_run_on_the_engine():
    """disparity_calculation.py:289,300-302,312: compute_disparity -> reprojectImageTo3D(filtered_img16, Q) -> mask."""
    from stereo_match_b200.synth import KITTI_Q
    l, r = synth_pair(320, 120, 64, seed=3)
    settings = sm.parse_config_file(None)
    settings.update(window_size=5, num_disparities=64)
    displ, filtered_img = sm.compute_disparity(l, r, settings)
    filtered_img16 = np.int16(filtered_img)
    points_3d = sm.reprojectImageTo3D(filtered_img16, KITTI_Q)
    mask = (filtered_img16 > 0) & (filtered_img16 < 1000)
    assert points_3d.shape == (120, 320, 3) and points_3d.dtype == np.float32 and mask.any()
    assert np.isfinite(points_3d[mask]).all()
