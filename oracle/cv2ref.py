"""Live cv2 4.13.0 -- the reference's real arithmetic (ORACLE: test infrastructure only).

The reference reaches these through stereo_vision/stereo_vision.py:153-166,178,209,
disparity_calculation.py:207-208,286-287,302 and rectified_img_cal.py:253-254.
"""
import numpy as np
import cv2

EXPECTED_VERSION = "4.13.0"


def version_ok():
    return cv2.__version__ == EXPECTED_VERSION


def sgbm_compute(left, right, **kw):
    m = cv2.StereoSGBM_create(**kw)
    return m.compute(np.ascontiguousarray(left), np.ascontiguousarray(right))


def bm_compute(left, right, numDisparities=64, blockSize=21, **setters):
    m = cv2.StereoBM_create(numDisparities=numDisparities, blockSize=blockSize)
    for k, v in setters.items():
        getattr(m, "set" + k[0].upper() + k[1:])(v)
    return m.compute(np.ascontiguousarray(left), np.ascontiguousarray(right))


def median3(img):
    return cv2.medianBlur(np.ascontiguousarray(img, np.int16), 3)


def filter_speckles(img, newVal, maxSize, maxDiff):
    out = np.array(img, np.int16, order="C", copy=True)
    cv2.filterSpeckles(out, newVal, maxSize, maxDiff)
    return out


def reproject(disp, Q, handleMissingValues=False):
    return cv2.reprojectImageTo3D(np.ascontiguousarray(disp), np.asarray(Q), handleMissingValues=handleMissingValues)


def remap(src, mapx, mapy):
    return cv2.remap(src, mapx, mapy, cv2.INTER_LINEAR)


def bgr2gray(bgr):
    return cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)


def kitti_rect_maps(w=1242, h=375):
    """cfg5 rectification maps, SURVEY.md 8(d) (shared input generator)."""
    from stereo_match_b200.synth import kitti_rect_maps as f
    return f(w, h)
