"""Seeded synthetic rectified stereo pairs (SURVEY.md 8(d) generator).

Input generation only -- never on a timed or parity-judged path.  Uses cv2 purely as an
image library (cubic resize / remap) so that bench inputs and test inputs are the very
pairs the survey's CPU numbers were quoted on.
"""
import numpy as np

KITTI_Q = np.array([[1, 0, 0, -609.5593],
                    [0, 1, 0, -172.854],
                    [0, 0, 0, 721.5377],
                    [0, 0, 1 / 0.54, 0]], np.float64)


def synth_pair(w, h, maxd, seed, noise=2.0):
    """Returns (left, right) uint8 (h, w): textured scene, ramp-plus-boxes disparity field."""
    import cv2
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (h // 4 + 2, (w + maxd) // 4 + 2), dtype=np.uint8)
    big = cv2.resize(base, (w + maxd, h), interpolation=cv2.INTER_CUBIC)
    fine = rng.integers(0, 256, (h, w + maxd), dtype=np.uint8)
    big = cv2.addWeighted(big, 0.7, fine, 0.3, 0)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    disp = (0.15 * maxd + 0.5 * maxd * yy / h).astype(np.float32)
    for _ in range(6):
        bx = int(rng.integers(0, max(w - w // 6, 1)))
        by = int(rng.integers(0, max(h - h // 6, 1)))
        disp[by:by + h // 6, bx:bx + w // 6] = np.float32(rng.uniform(0.1, 0.9) * maxd)
    left = np.ascontiguousarray(big[:, maxd:maxd + w])
    # right(x) = left(x + d): a scene point at left column x appears at right column x - d
    right = cv2.remap(big, xx + maxd + disp, yy, cv2.INTER_LINEAR).astype(np.float32)
    right += rng.normal(0.0, noise, right.shape).astype(np.float32)
    right = np.clip(np.rint(right), 0, 255).astype(np.uint8)
    return left, right


def kitti_rect_maps(w=1242, h=375):
    """Rectification maps of cfg5: float32 mapx, mapy from a mild distortion
    plus a small rotation.  Input generation only (cv2 used as a calibration library)."""
    import cv2
    K = np.array([[721.5377, 0, 609.5593], [0, 721.5377, 172.854], [0, 0, 1]], np.float64)
    dist = np.array([-0.05, 0.01, 5e-4, -3e-4, 0], np.float64)
    R, _ = cv2.Rodrigues(np.array([0.002, -0.004, 0.003], np.float64))
    return cv2.initUndistortRectifyMap(K, dist, R, K, (w, h), cv2.CV_32FC1)
