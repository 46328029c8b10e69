"""ctypes binding of oracle/stereo_oracle.c (ORACLE: test infrastructure only)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_SRC = os.path.join(_HERE, "stereo_oracle.c")
_SRCS = [_SRC, os.path.join(_HERE, "wls_oracle.c")]


def build(force=False):
    """Compile the C restatement with gcc (seconds)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in _SRCS):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-fvisibility=hidden",
                               "-o", _SO] + _SRCS + ["-lm"])
    return _SO


class SgbmParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
        "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


class BmParams(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "numDisparities", "blockSize", "preFilterCap", "textureThreshold", "uniquenessRatio", "minDisparity")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def sgbm_params(minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0,
                preFilterCap=0, uniquenessRatio=0, speckleWindowSize=0, speckleRange=0, mode=0):
    return SgbmParams(minDisparity, numDisparities, blockSize, P1, P2, disp12MaxDiff,
                      preFilterCap, uniquenessRatio, speckleWindowSize, speckleRange, mode)


def sgbm_geometry(W, **kw):
    out = (C.c_int * 8)()
    p = sgbm_params(**kw)
    lib().orc_sgbm_geometry(C.byref(p), int(W), out)
    keys = ("minX1", "maxX1", "width1", "INVALID", "P1", "P2", "ftzero", "uniq")
    return dict(zip(keys, list(out)))


def sgbm_compute(left, right, want_volumes=False, **kw):
    """Returns disp (H,W) int16; with want_volumes also (C, S, raw).  Images: (H,W) gray or (H,W,3)."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    H, W = left.shape[:2]
    cn = 1 if left.ndim == 2 else left.shape[2]
    p = sgbm_params(**kw)
    disp = np.empty((H, W), np.int16)
    Cv = Sv = raw = None
    if want_volumes:
        g = sgbm_geometry(W, **kw)
        w1 = max(g["width1"], 0)
        Cv = np.zeros((H, w1, p.numDisparities), np.int16)
        Sv = np.zeros_like(Cv)
        raw = np.empty((H, W), np.int16)
    rc = lib().orc_sgbm_compute_cn(_p(left), _p(right), H, W, cn, C.byref(p), _p(disp), _p(Cv), _p(Sv), _p(raw))
    if rc == -2:
        raise ValueError("oracle: cv2.error -- width - (minDisparity + numDisparities) > SADWindowSize/2")
    if rc != 0:
        raise ValueError("oracle: unsupported SGBM mode / channel count / tap")
    return (disp, Cv, Sv, raw) if want_volumes else disp


def sgbm_hsum(left, right, **kw):
    """Stage tap: horizontal window sums hs[H][width1][D] int32."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    H, W = left.shape
    p = sgbm_params(**kw)
    w1 = sgbm_geometry(W, **kw)["width1"]
    out = np.empty((H, w1, p.numDisparities), np.int32)
    if lib().orc_sgbm_hsum(_p(left), _p(right), H, W, C.byref(p), _p(out)) != 0:
        raise ValueError("oracle: empty computed area")
    return out


def sgbm_finish(S, W, **kw):
    """Stage tap: S volume -> final disparity (WTA, LR check, median, speckle filter)."""
    S = np.ascontiguousarray(S, np.int16)
    p = sgbm_params(**kw)
    disp = np.empty((S.shape[0], W), np.int16)
    lib().orc_sgbm_finish(_p(S), S.shape[0], int(W), C.byref(p), _p(disp))
    return disp


def sgbm_wta(S, W, **kw):
    """Stage tap: S volume -> raw disparity rows (before the median)."""
    S = np.ascontiguousarray(S, np.int16)
    p = sgbm_params(**kw)
    raw = np.empty((S.shape[0], W), np.int16)
    lib().orc_sgbm_wta(_p(S), S.shape[0], int(W), C.byref(p), _p(raw))
    return raw


def sgbm_path(Cvol, P1, P2, k):
    """L_r of direction k (0..7, order of DIRS in stereo_oracle.c) from an int16 C volume."""
    Cvol = np.ascontiguousarray(Cvol, np.int16)
    H, w1, D = Cvol.shape
    out = np.empty_like(Cvol)
    lib().orc_sgbm_path(_p(Cvol), H, w1, D, int(P1), int(P2), int(k), _p(out))
    return out


def bm_compute(left, right, numDisparities=64, blockSize=21, preFilterCap=31,
               textureThreshold=10, uniquenessRatio=15, speckleWindowSize=0, speckleRange=0, minDisparity=0,
               disp12MaxDiff=-1):
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    H, W = left.shape
    if 0 <= disp12MaxDiff < 100000:
        raise ValueError("oracle: StereoBM's left-right check is only restated for disp12MaxDiff < 0 (off) or so large "
                         "that it never rejects (the ximgproc factories set 1000000)")
    p = BmParams(numDisparities, blockSize, preFilterCap, textureThreshold, uniquenessRatio, minDisparity)
    disp = np.empty((H, W), np.int16)
    rc = lib().orc_bm_compute(_p(left), _p(right), H, W, C.byref(p), _p(disp))
    if rc != 0:
        raise ValueError("oracle: bad BM parameters")
    # cv2 4.13.0 StereoBM: the speckle range is passed to filterSpeckles unscaled (StereoSGBM passes 16 * range) and
    # a negative range skips the filter (probed: tests/test_oracle_vs_cv2.py)
    if speckleWindowSize > 0 and speckleRange >= 0:
        disp = filter_speckles(disp, (minDisparity - 1) * 16, speckleWindowSize, speckleRange)
    return disp


def median3(img):
    img = np.ascontiguousarray(img, np.int16)
    out = np.empty_like(img)
    lib().orc_median3_i16(_p(img), img.shape[0], img.shape[1], _p(out))
    return out


def filter_speckles(img, newVal, maxSize, maxDiff):
    out = np.array(img, np.int16, order="C", copy=True)
    lib().orc_filter_speckles(_p(out), out.shape[0], out.shape[1], int(newVal), int(maxSize), int(maxDiff))
    return out


def reproject(disp, Q, handleMissingValues=False):
    d = np.ascontiguousarray(disp).astype(np.float32)      # no /16: SURVEY.md A.6
    Q = np.ascontiguousarray(Q, np.float64)
    H, W = d.shape
    out = np.empty((H, W, 3), np.float32)
    lib().orc_reproject(_p(d), H, W, _p(Q), int(bool(handleMissingValues)), _p(out))
    return out


def remap(src, mapx, mapy):
    src = np.ascontiguousarray(src, np.uint8)
    cn = 1 if src.ndim == 2 else src.shape[2]
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    H, W = mapx.shape
    out = np.empty((H, W) if src.ndim == 2 else (H, W, cn), np.uint8)
    lib().orc_remap_u8(_p(src), src.shape[0], src.shape[1], cn, _p(mapx), _p(mapy), H, W, _p(out))
    return out


def bgr2gray(bgr):
    bgr = np.ascontiguousarray(bgr, np.uint8)
    out = np.empty(bgr.shape[:2], np.uint8)
    lib().orc_bgr2gray(_p(bgr), C.c_size_t(out.size), _p(out))
    return out


class WlsParams(C.Structure):
    _fields_ = [("roi_x", C.c_int), ("roi_y", C.c_int), ("roi_w", C.c_int), ("roi_h", C.c_int), ("min_disp", C.c_int),
                ("use_confidence", C.c_int), ("lrc_thresh", C.c_int), ("radius", C.c_int), ("roll_off", C.c_float),
                ("lambda_", C.c_float), ("sigma", C.c_float), ("num_iter", C.c_int), ("attenuation", C.c_float)]


def wls_filter(disp_left, disp_right, guide, roi, min_disp=0, use_confidence=True, lambda_=8000.0, sigma=1.0,
               lrc_thresh=24, radius=5, roll_off=0.001, num_iter=3, attenuation=0.25):
    """oracle/wls_oracle.c (PARITY UNPINNED restatement of cv2.ximgproc's DisparityWLSFilter) -> (filtered int16,
    confidence float32)."""
    dl = np.ascontiguousarray(disp_left, np.int16)
    dr = np.ascontiguousarray(disp_right, np.int16) if disp_right is not None else None
    g = np.ascontiguousarray(guide, np.uint8)
    H, W = dl.shape
    cn = 1 if g.ndim == 2 else g.shape[2]
    p = WlsParams(int(roi[0]), int(roi[1]), int(roi[2]), int(roi[3]), int(min_disp), int(bool(use_confidence)),
                  int(lrc_thresh), int(radius), float(roll_off), float(lambda_), float(sigma), int(num_iter),
                  float(attenuation))
    out = np.empty((H, W), np.int16)
    conf = np.empty((H, W), np.float32)
    rc = lib().orc_wls_filter(_p(dl), _p(dr), _p(g), cn, H, W, C.byref(p), _p(out), _p(conf))
    if rc != 0:
        raise ValueError("oracle: bad WLS arguments")
    return out, conf
