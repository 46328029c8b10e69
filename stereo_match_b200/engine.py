"""cv2-shaped Python surface over libb2stereo.so.

Mirrors what the reference calls on its hot path:
  cv2.StereoSGBM_create / .compute     stereo_vision/stereo_vision.py:153-163,178
  cv2.StereoBM_create / .compute       stereo_vision/stereo_vision.py:165-166,178
  cv2.reprojectImageTo3D               stereo_vision/stereo_vision.py:209, disparity_calculation.py:302
  cv2.remap(..., INTER_LINEAR)         rectified_img_cal.py:253-254, disparity_calculation.py:207-208
  cv2.cvtColor(..., COLOR_BGR2GRAY)    disparity_calculation.py:286-287

Arrays may be numpy (copied host->device and back), torch CUDA tensors or DLPack producers
(zero copy); the result is returned as the same kind.  A leading batch axis is accepted
everywhere ([B,H,W]) -- that is how pairs are kept in flight on one GPU.
PyTorch is used for device memory and streams only; all arithmetic is in the CUDA library.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import error

STEREO_SGBM_MODE_SGBM = 0
STEREO_SGBM_MODE_HH = 1
STEREO_SGBM_MODE_SGBM_3WAY = 2
STEREO_SGBM_MODE_HH4 = 3
INTER_LINEAR = 1
COLOR_BGR2GRAY = 6
CV_32F = 5


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("stereo_match_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")


def _to_device(a, name="array"):
    """-> (contiguous torch CUDA tensor, kind) where kind is 'numpy' or 'torch'."""
    _require_cuda()
    if isinstance(a, np.ndarray):
        return torch.from_numpy(np.ascontiguousarray(a)).cuda(non_blocking=False), "numpy"
    if isinstance(a, torch.Tensor):
        if not a.is_cuda:
            return a.contiguous().cuda(), "torch_cpu"
        return a.contiguous(), "torch"
    if hasattr(a, "__dlpack__"):
        t = torch.from_dlpack(a)
        if not t.is_cuda:
            t = t.cuda()
        return t.contiguous(), "torch"
    raise error("%s: expected a numpy array, torch tensor or DLPack producer, got %r" % (name, type(a)))


def _from_device(t, kind):
    if kind == "numpy":
        return t.cpu().numpy()
    if kind == "torch_cpu":
        return t.cpu()
    return t


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


class _Matcher:
    def __init__(self):
        self._h = C.c_void_p()

    def __del__(self):
        try:
            if self._h:
                _lib.load().b2s_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _check_pair(self, who, left, right):
        """-> (left, right) as [B, H, W] or [B, H, W, 3] uint8 device tensors, kind, squeeze, channels.
        Shapes: (H, W) gray, (H, W, 3) colour (what cv2 takes), (B, H, W) gray batch, (B, H, W, 3) colour batch."""
        l, kind = _to_device(left, "left")
        r, _ = _to_device(right, "right")
        if l.shape != r.shape or l.dtype != r.dtype or l.dtype != torch.uint8:
            raise error("%s: left.size() == right.size() && left.type() == right.type() && left.depth() == CV_8U" % who)
        cn = 1
        if l.dim() == 4 and l.shape[3] == 3:
            cn, squeeze = 3, False
        elif l.dim() == 3 and l.shape[2] == 3 and l.shape[1] != 3:
            l, r, cn, squeeze = l[None], r[None], 3, True
        elif l.dim() == 2:
            l, r, squeeze = l[None], r[None], True
        elif l.dim() == 3:
            squeeze = False
        else:
            raise error("%s: images must be (H, W), (H, W, 3), (B, H, W) or (B, H, W, 3) uint8" % who)
        return l.contiguous(), r.contiguous(), kind, squeeze, cn


class StereoSGBM(_Matcher):
    """Drop-in for the object cv2.StereoSGBM_create returns."""

    _FIELDS = ("minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
               "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")

    def __init__(self, **kw):
        super().__init__()
        self._p = _lib.SgbmParams(**{k: int(kw[k]) for k in self._FIELDS})
        _lib.check(_lib.load().b2s_sgbm_create(C.byref(self._p), C.byref(self._h)))

    def _push(self):
        _lib.check(_lib.load().b2s_sgbm_set_params(self._h, C.byref(self._p)))

    def compute(self, left, right, disparity=None):
        l, r, kind, squeeze, cn = self._check_pair("StereoSGBM", left, right)
        B, H, W = l.shape[:3]
        out = torch.empty((B, H, W), dtype=torch.int16, device=l.device)
        with torch.cuda.device(l.device):
            _lib.check(_lib.load().b2s_sgbm_compute_cn(self._h, _ptr(l), _ptr(r), H, W, cn, B, _ptr(out), _stream()))
        if squeeze:
            out = out[0]
        res = _from_device(out, kind)
        if disparity is not None:
            if isinstance(disparity, np.ndarray):
                disparity[...] = res if isinstance(res, np.ndarray) else res.cpu().numpy()
            else:
                disparity.copy_(out)
            return disparity
        return res

    def geometry(self, W):
        g = _lib.SgbmGeometry()
        _lib.check(_lib.load().b2s_sgbm_get_geometry(self._h, int(W), C.byref(g)))
        return {n: getattr(g, n) for n, _ in g._fields_}

    def workspace_bytes(self, H, W, batch=1):
        return int(_lib.load().b2s_sgbm_workspace_bytes(self._h, int(H), int(W), int(batch)))

    def overflowed(self):
        """True if the last compute left the bit-exact safe domain."""
        f = C.c_int(0)
        _lib.check(_lib.load().b2s_sgbm_overflow_flag(self._h, C.byref(f)))
        return bool(f.value)

    STAGES = ("descriptors", "cost", "aggregate", "wta_lr", "median", "speckle")

    def enable_timing(self, on=True):
        _lib.check(_lib.load().b2s_sgbm_enable_timing(self._h, int(bool(on))))

    def stage_ms(self):
        """Per-stage device time of the last compute (CUDA events on its stream)."""
        ms = (C.c_float * len(self.STAGES))()
        _lib.check(_lib.load().b2s_sgbm_stage_ms(self._h, ms))
        return dict(zip(self.STAGES, [float(v) for v in ms]))

    def sweep_ms(self):
        """(forward sweep, mirrored sweep + winner-take-all) of the last timed compute, in milliseconds."""
        ms = (C.c_float * 2)()
        _lib.check(_lib.load().b2s_sgbm_sweep_ms(self._h, ms))
        return float(ms[0]), float(ms[1])

    def keep_volumes(self, on=True):
        """Debug tap: make the following computes also write the full S volume."""
        _lib.check(_lib.load().b2s_sgbm_keep_volumes(self._h, int(bool(on))))

    def debug_volumes(self, B, H, W):
        """(cost, S, geometry) of the last compute; torch int16 views of the uint16 volumes
        [B, H, width1, Dp] (test tap; needs keep_volumes() before the compute)."""
        g = self.geometry(W)
        n = B * H * g["width1"] * g["Dp"]
        cost = torch.empty(n, dtype=torch.int16, device="cuda")
        S = torch.empty(n, dtype=torch.int16, device="cuda")
        _lib.check(_lib.load().b2s_sgbm_debug_volumes(self._h, _ptr(cost), _ptr(S), n, _stream()))
        shape = (B, H, g["width1"], g["Dp"])
        cost, S = cost.view(shape), S.view(shape)
        if g["swizzled"]:                 # records are stored in the wavefront kernel's lane order: undo it
            d = torch.arange(g["Dp"], device="cuda")
            j, k = d // 16, d % 16        # b2s_cell_pos (common.cuh): register i packs cells (16j + i, 16j + 8 + i)
            h, i = k // 8, k % 8
            pos = (((i // 4) * g["G"] + j) * 4 + i % 4) * 2 + h
            cost, S = cost[..., pos], S[..., pos]
        return cost, S, g


def _make_accessors(cls, fields):
    for f in fields:
        cap = f[0].upper() + f[1:]

        def getter(self, _f=f):
            return int(getattr(self._p, _f))

        def setter(self, v, _f=f):
            setattr(self._p, _f, int(v))
            self._push()

        setattr(cls, "get" + cap, getter)
        setattr(cls, "set" + cap, setter)


_make_accessors(StereoSGBM, StereoSGBM._FIELDS)


def StereoSGBM_create(minDisparity=0, numDisparities=16, blockSize=3, P1=0, P2=0, disp12MaxDiff=0,
                      preFilterCap=0, uniquenessRatio=0, speckleWindowSize=0, speckleRange=0,
                      mode=STEREO_SGBM_MODE_SGBM):
    """Same signature and defaults as cv2.StereoSGBM_create."""
    return StereoSGBM(minDisparity=minDisparity, numDisparities=numDisparities, blockSize=blockSize, P1=P1, P2=P2,
                      disp12MaxDiff=disp12MaxDiff, preFilterCap=preFilterCap, uniquenessRatio=uniquenessRatio,
                      speckleWindowSize=speckleWindowSize, speckleRange=speckleRange, mode=mode)


class StereoBM(_Matcher):
    """Drop-in for cv2.StereoBM_create(numDisparities, blockSize) (XSOBEL prefilter; any minDisparity; disp12MaxDiff
    either off (< 0, the cv2 default) or so large that it never rejects, as cv2.ximgproc's factories set it)."""

    _FIELDS = ("numDisparities", "blockSize", "preFilterCap", "textureThreshold", "uniquenessRatio",
               "speckleWindowSize", "speckleRange", "minDisparity", "disp12MaxDiff")

    def __init__(self, numDisparities=0, blockSize=21):
        super().__init__()
        # cv2.StereoBM_create: any numDisparities <= 0 means 64 (probed: -16, -1, 0 all give getNumDisparities() == 64)
        self._p = _lib.BmParams(numDisparities=int(numDisparities) if int(numDisparities) > 0 else 64, blockSize=int(blockSize),
                                preFilterCap=31, textureThreshold=10, uniquenessRatio=15,
                                speckleWindowSize=0, speckleRange=0, minDisparity=0, disp12MaxDiff=-1)
        _lib.check(_lib.load().b2s_bm_create(C.byref(self._p), C.byref(self._h)))

    def _push(self):
        # BM parameters are read at compute time from a fresh handle
        lib = _lib.load()
        lib.b2s_destroy(self._h)
        self._h = C.c_void_p()
        _lib.check(lib.b2s_bm_create(C.byref(self._p), C.byref(self._h)))

    def compute(self, left, right, disparity=None):
        l, kind = _to_device(left, "left")
        r, _ = _to_device(right, "right")
        if l.dtype != torch.uint8 or r.dtype != torch.uint8 or (l.dim() == 3 and l.shape[2] == 3 and l.shape[1] != 3):
            raise error("StereoBM: Both input images must have CV_8UC1")
        if l.shape != r.shape:
            raise error("StereoBM: All the images must have the same size")
        squeeze = l.dim() == 2
        if squeeze:
            l, r = l[None], r[None]
        B, H, W = l.shape
        out = torch.empty((B, H, W), dtype=torch.int16, device=l.device)
        with torch.cuda.device(l.device):
            _lib.check(_lib.load().b2s_bm_compute(self._h, _ptr(l), _ptr(r), H, W, B, _ptr(out), _stream()))
        if squeeze:
            out = out[0]
        res = _from_device(out, kind)
        if disparity is not None:
            if isinstance(disparity, np.ndarray):
                disparity[...] = res
            else:
                disparity.copy_(out)
            return disparity
        return res


_make_accessors(StereoBM, StereoBM._FIELDS)


def StereoBM_create(numDisparities=0, blockSize=21):
    return StereoBM(numDisparities=numDisparities, blockSize=blockSize)
