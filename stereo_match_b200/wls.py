"""cv2.ximgproc's disparity WLS filter as the reference uses it (stereo_vision/stereo_vision.py:171-182,
disparity_test.py:180-199, mapTo3D_mc_cnn.py:94-100): createDisparityWLSFilter(left_matcher),
createDisparityWLSFilterGeneric(use_confidence), .setLambda / .setSigmaColor, .filter(displ, gray_l, None, dispr).

PARITY UNPINNED.  cv2.ximgproc (opencv_contrib) exists neither in this image nor on the GPU box and the reference has
no fixture for it.  The factory rules below (offsets of the valid-disparity rectangle, the depth-discontinuity radius,
the parameters the factory overwrites on the left matcher) and the defaults are restated from the opencv_contrib 4.x
sources from memory; the CUDA kernels (csrc/wls.cu) are bit-exact with the restatement in oracle/wls_oracle.c.
"""
import ctypes as C
import math

import numpy as np
import torch

from . import _lib, engine
from ._lib import error
from .engine import _to_device, _from_device, _ptr, _stream


class DisparityWLSFilter:
    """Counterpart of cv2.ximgproc.DisparityWLSFilter (fast-global-smoother based, optional LR confidence)."""

    def __init__(self, use_confidence, left_offset=0, right_offset=0, top_offset=0, bottom_offset=0, min_disp=0):
        self._h = C.c_void_p()
        _lib.check(_lib.load().b2s_wls_create(C.byref(self._h)))
        self._use_confidence = bool(use_confidence)
        self._offsets = (int(left_offset), int(right_offset), int(top_offset), int(bottom_offset))
        self._min_disp = int(min_disp)
        self._lambda = 8000.0
        self._sigma = 1.0
        self._lrc_thresh = 24
        self._radius = 5
        self._roll_off = 0.001
        self._conf = None
        self._roi = (0, 0, 0, 0)

    def __del__(self):
        try:
            if self._h:
                _lib.load().b2s_wls_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # ---- cv2.ximgproc.DisparityWLSFilter accessors -------------------------------------------------------------
    def getLambda(self):
        return self._lambda

    def setLambda(self, v):
        self._lambda = float(v)

    def getSigmaColor(self):
        return self._sigma

    def setSigmaColor(self, v):
        self._sigma = float(v)

    def getLRCthresh(self):
        return self._lrc_thresh

    def setLRCthresh(self, v):
        self._lrc_thresh = int(v)

    def getDepthDiscontinuityRadius(self):
        return self._radius

    def setDepthDiscontinuityRadius(self, v):
        self._radius = int(v)

    def getConfidenceMap(self):
        """float32 (H, W) confidence of the last filter() call (0..255), like cv2."""
        return self._conf

    def getROI(self):
        """(x, y, w, h) of the valid-disparity rectangle used by the last filter() call."""
        return self._roi

    def filter(self, disparity_map_left, left_view, filtered_disparity_map=None, disparity_map_right=None, ROI=None,
               right_view=None):
        """filter(displ, gray_l, None, dispr) of stereo_vision.py:182 -> int16 filtered disparity (same array kind
        as `disparity_map_left`).  Batched (B, H, W) disparities with (B, H, W) / (B, H, W, 3) views are accepted."""
        dl, kind = _to_device(disparity_map_left, "disparity_map_left")
        g, _ = _to_device(left_view, "left_view")
        if dl.dtype != torch.int16:
            raise error("DisparityWLSFilter: disparity_map_left.depth() == CV_16S && disparity_map_left.channels() == 1")
        if g.dtype != torch.uint8:
            raise error("DisparityWLSFilter: left_view.depth() == CV_8U && (left_view.channels() == 3 || left_view.channels() == 1)")
        dr = None
        if disparity_map_right is not None:
            dr, _ = _to_device(disparity_map_right, "disparity_map_right")
            if dr.dtype != torch.int16 or dr.shape != dl.shape:
                raise error("DisparityWLSFilter: disparity_map_right must be CV_16S of the size of disparity_map_left")
        elif self._use_confidence:
            raise error("DisparityWLSFilter: !disparity_map_right.empty() (the confidence map needs the right-view disparity)")
        squeeze = dl.dim() == 2
        if squeeze:
            dl = dl[None]
            dr = dr[None] if dr is not None else None
            g = g[None]
        if dl.dim() != 3:
            raise error("DisparityWLSFilter: disparity_map_left must be (H, W) or (B, H, W)")
        B, H, W = dl.shape
        cn = 3 if (g.dim() == 4 and g.shape[3] == 3) else 1
        if tuple(g.shape[:3]) != (B, H, W) or (g.dim() == 4 and g.shape[3] not in (1, 3)) or g.dim() not in (3, 4):
            raise error("DisparityWLSFilter: left_view must have the size of the disparity map and 1 or 3 channels")
        lo, ro, to, bo = self._offsets
        if ROI is not None and ROI[2] * ROI[3] != 0:
            roi = tuple(int(v) for v in ROI)
        else:
            roi = (lo, to, max(W - lo - ro, 0), max(H - to - bo, 0))
        self._roi = roi
        p = _lib.WlsParams(roi_x=roi[0], roi_y=roi[1], roi_w=roi[2], roi_h=roi[3], min_disp=self._min_disp,
                           use_confidence=int(self._use_confidence), lrc_thresh=self._lrc_thresh, radius=self._radius,
                           roll_off=self._roll_off, lambda_=self._lambda, sigma=self._sigma, num_iter=3, attenuation=0.25)
        out = torch.empty((B, H, W), dtype=torch.int16, device=dl.device)
        conf = torch.empty((B, H, W), dtype=torch.float32, device=dl.device) if self._use_confidence else None
        with torch.cuda.device(dl.device):
            _lib.check(_lib.load().b2s_wls_filter(self._h, _ptr(dl), _ptr(dr) if dr is not None else None, _ptr(g), cn,
                                                  H, W, B, C.byref(p), _ptr(out), _ptr(conf) if conf is not None else None,
                                                  _stream()))
        if conf is not None:
            self._conf = _from_device(conf[0] if squeeze else conf, kind)
        res = _from_device(out[0] if squeeze else out, kind)
        if filtered_disparity_map is not None:
            if isinstance(filtered_disparity_map, np.ndarray):
                filtered_disparity_map[...] = res
            else:
                filtered_disparity_map.copy_(out[0] if squeeze else out)
            return filtered_disparity_map
        return res


def createDisparityWLSFilterGeneric(use_confidence):
    """cv2.ximgproc.createDisparityWLSFilterGeneric (disparity_test.py:182): no matcher-specific rectangle."""
    return DisparityWLSFilter(use_confidence)


def createDisparityWLSFilter(matcher_left):
    """cv2.ximgproc.createDisparityWLSFilter(left_matcher) (stereo_vision.py:172).

    Like the original it RE-PARAMETRISES the left matcher so that its raw map suits the filter -- disp12MaxDiff =
    1000000, speckleWindowSize = 0, uniquenessRatio = 0 (StereoBM also textureThreshold = 0) -- and derives the
    valid-disparity rectangle and the depth-discontinuity radius from it."""
    if isinstance(matcher_left, engine.StereoBM):
        min_disp = matcher_left.getMinDisparity()
        num_disp, wsize = matcher_left.getNumDisparities(), matcher_left.getBlockSize()
        matcher_left.setDisp12MaxDiff(1000000)
        matcher_left.setSpeckleWindowSize(0)
        matcher_left.setTextureThreshold(0)
        matcher_left.setUniquenessRatio(0)
        w2 = wsize // 2
        f = DisparityWLSFilter(True, max(0, min_disp + num_disp) + w2, max(0, -min_disp) + w2, w2, w2, min_disp)
        f.setDepthDiscontinuityRadius(int(math.ceil(0.33 * wsize)))
        return f
    if isinstance(matcher_left, engine.StereoSGBM):
        min_disp = matcher_left.getMinDisparity()
        num_disp, wsize = matcher_left.getNumDisparities(), matcher_left.getBlockSize()
        matcher_left.setDisp12MaxDiff(1000000)
        matcher_left.setSpeckleWindowSize(0)
        matcher_left.setUniquenessRatio(0)
        f = DisparityWLSFilter(True, max(0, min_disp + num_disp), max(0, -min_disp), 0, 0, min_disp)
        f.setDepthDiscontinuityRadius(int(math.ceil(0.5 * wsize)))
        return f
    raise error("DisparityWLSFilter natively supports only StereoBM and StereoSGBM")
