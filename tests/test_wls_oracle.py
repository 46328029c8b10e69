"""CPU checks of oracle/wls_oracle.c, the restatement of cv2.ximgproc's DisparityWLSFilter (reference call sites
stereo_vision/stereo_vision.py:171-182, disparity_test.py:180-199).

PARITY UNPINNED: there is no cv2.ximgproc in this image (nor on the GPU box) and the reference holds no fixture, so
these tests pin the restatement to the PROPERTIES of the published algorithm, not to cv2 outputs."""
import numpy as np
import pytest


def _maps(h=40, w=70, D=16, seed=0):
    rng = np.random.default_rng(seed)
    g = rng.integers(0, 256, (h, w), dtype=np.uint8)
    dl = (rng.integers(0, D, (h, w)) * 16).astype(np.int16)
    dr = (-dl).astype(np.int16)
    return g, dl, dr


def test_no_cv2_ximgproc_here():
    """The reason for 'unpinned': if this ever fails, pin the restatement against cv2.ximgproc instead."""
    cv2 = pytest.importorskip("cv2")
    assert not hasattr(cv2, "ximgproc")


def test_constant_map_is_a_fixed_point(port):
    g, _, _ = _maps()
    d = np.full(g.shape, 320, np.int16)
    out, conf = port.wls_filter(d, (-d).astype(np.int16), g, (16, 0, 54, 40), lambda_=80000, sigma=1.2, radius=3)
    assert np.all(out[:, 16:] == 320) and np.all(out[:, :16] == -16)          # outside the rectangle: 16 * (minD - 1)
    assert np.all(conf[:, 16:] == 255.0) and np.all(conf[:, :16] == 0.0)


def test_lambda_zero_returns_the_input(port):
    g, dl, dr = _maps(seed=1)
    out, _ = port.wls_filter(dl, None, g, (0, 0, 70, 40), use_confidence=False, lambda_=0.0)
    assert np.array_equal(out, dl)


def test_smoothing_preserves_the_mean_and_reduces_variation(port):
    g, dl, _ = _maps(seed=2)
    g[:] = 128                                     # flat guide: plain isotropic smoothing
    out, _ = port.wls_filter(dl, None, g, (0, 0, 70, 40), use_confidence=False, lambda_=50.0, sigma=1.5)
    assert abs(float(out.mean()) - float(dl.mean())) < 1.0       # (I + lambda A) has unit row sums: the mean survives
    assert out.std() < 0.5 * dl.std()


def test_guide_edges_stop_the_smoothing(port):
    h, w = 30, 60
    g = np.zeros((h, w), np.uint8)
    g[:, 30:] = 255
    d = np.zeros((h, w), np.int16)
    d[:, 30:] = 800
    out, _ = port.wls_filter(d, None, g, (0, 0, w, h), use_confidence=False, lambda_=8000.0, sigma=1.0)
    assert np.abs(out.astype(int) - d).max() <= 1                # exp(-255 / sigma) ~ 0: the two halves do not mix
    g[:] = 7
    out2, _ = port.wls_filter(d, None, g, (0, 0, w, h), use_confidence=False, lambda_=8000.0, sigma=1.0)
    assert 300 < out2[:, 29].mean() < 500 and 300 < out2[:, 30].mean() < 500     # no edge in the guide: they do


def test_inconsistent_pixels_lose_their_vote(port):
    """LR check: where left and right disagree the confidence is 0 and the value is filled in from the neighbours."""
    g, _, _ = _maps(seed=3)
    g[:] = 100
    d = np.full(g.shape, 160, np.int16)
    dr = np.full(g.shape, -160, np.int16)
    d[10:14, 40:44] = 640                                         # a blob the right view does not confirm
    out, conf = port.wls_filter(d, dr, g, (16, 0, 54, 40), lambda_=8000.0, sigma=1.5, radius=1)
    assert np.all(conf[10:14, 40:44] == 0.0)
    assert np.abs(out[10:14, 40:44].astype(int) - 160).max() <= 2
    assert np.abs(out[30:, 20:30].astype(int) - 160).max() == 0


def test_three_channel_guide_and_bad_arguments(port):
    g, dl, dr = _maps(seed=4)
    g3 = np.stack([g, g, g], -1)
    a, _ = port.wls_filter(dl, dr, g, (16, 0, 54, 40), lambda_=800.0, sigma=4.0)
    b, _ = port.wls_filter(dl, dr, g3, (16, 0, 54, 40), lambda_=800.0, sigma=4.0 * np.sqrt(3.0))
    assert np.abs(a.astype(int) - b).max() <= 1                   # |dI| of three equal channels = sqrt(3) |dI|
    with pytest.raises(ValueError):
        port.wls_filter(dl, None, g, (16, 0, 54, 40))             # confidence needs the right-view map
    with pytest.raises(ValueError):
        port.wls_filter(dl, dr, g, (16, 0, 60, 40))               # rectangle leaves the image
