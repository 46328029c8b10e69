"""Stand-alone stages with cv2 signatures: reprojectImageTo3D, remap, cvtColor(BGR2GRAY),
medianBlur(ksize=3, int16) and filterSpeckles.  See engine.py for the array conventions."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import error
from .engine import _to_device, _from_device, _ptr, _stream, INTER_LINEAR, COLOR_BGR2GRAY, CV_32F

_DT = {torch.uint8: _lib.DT_U8, torch.int16: _lib.DT_I16, torch.int32: _lib.DT_I32, torch.float32: _lib.DT_F32}


def reprojectImageTo3D(disparity, Q, _3dImage=None, handleMissingValues=False, ddepth=-1, *, disp_scale=1.0):
    """cv2.reprojectImageTo3D (stereo_vision.py:209, disparity_calculation.py:302).

    u8/int16/int32/float32 disparity is converted to float32 WITHOUT scaling, like cv2.
    `disp_scale` (keyword only, extension) multiplies it on the device first, which fuses
    project_points_3D's "/16.0".  Returns (H, W, 3) float32 ([B,H,W,3] for batched input).
    """
    if ddepth not in (-1, CV_32F):
        raise error("reprojectImageTo3D: only ddepth=-1/CV_32F is supported", _lib.E_UNSUPPORTED)
    d, kind = _to_device(disparity, "disparity")
    if d.dtype not in _DT:
        raise error("reprojectImageTo3D: stype == CV_8UC1 || stype == CV_16SC1 || stype == CV_32SC1 || stype == CV_32FC1")
    Qh = np.ascontiguousarray(Q.detach().cpu().numpy() if isinstance(Q, torch.Tensor) else Q, dtype=np.float64)
    if Qh.shape != (4, 4):
        raise error("reprojectImageTo3D: Q.size() == Size(4,4)")
    squeeze = d.dim() == 2
    if squeeze:
        d = d[None]
    if d.dim() != 3:
        raise error("reprojectImageTo3D: disparity must be (H, W) or (B, H, W)")
    B, H, W = d.shape
    out = torch.empty((B, H, W, 3), dtype=torch.float32, device=d.device)
    scratch = torch.empty(B, dtype=torch.float32, device=d.device) if handleMissingValues else None
    with torch.cuda.device(d.device):
        _lib.check(_lib.load().b2s_reproject(
            _ptr(d), _DT[d.dtype], H, W, B, Qh.ctypes.data_as(C.POINTER(C.c_double)), C.c_float(disp_scale),
            int(bool(handleMissingValues)), _ptr(scratch) if scratch is not None else None, _ptr(out), _stream()))
    if squeeze:
        out = out[0]
    res = _from_device(out, kind)
    if _3dImage is not None:
        if isinstance(_3dImage, np.ndarray):
            _3dImage[...] = res
        else:
            _3dImage.copy_(out)
        return _3dImage
    return res


_DT_ANY = dict(_DT)
_DT_ANY.update({torch.uint16: _lib.DT_U16, torch.float64: _lib.DT_F64})


def map2Dto3D(pt2D, ptd, cx, cy, fx, fy, threshold):
    """map2Dto3D of the reference (mapTo3D.py:35-56, called at :116): 3-D points of a depth image,
    u = -(col - cx) * z / fx, v = (row - cy) * z / fy, z, in float64 like the reference's numpy expression.

    pt2D (H, W) selects the pixels (pt2D > threshold), ptd (H, W) is the depth map.  Returns (H, W, 3) float64.
    Like the reference, which reshapes the masked result back to the image shape, it raises ValueError unless
    every pixel passes the threshold.  Batched (B, H, W) input is accepted as an extension."""
    a, kind = _to_device(pt2D, "pt2D")
    z, _ = _to_device(ptd, "ptd")
    if a.dtype == torch.bool:
        a = a.to(torch.uint8)
    if a.dtype not in _DT_ANY or z.dtype not in _DT_ANY:
        raise error("map2Dto3D: images must be uint8, uint16, int16, int32, float32 or float64")
    if a.shape != z.shape or a.dim() not in (2, 3):
        raise error("map2Dto3D: pt2D and ptd must be (H, W) arrays of the same shape")
    squeeze = a.dim() == 2
    if squeeze:
        a, z = a[None], z[None]
    B, H, W = a.shape
    out = torch.empty((B, H, W, 3), dtype=torch.float64, device=a.device)
    npass = torch.zeros(1, dtype=torch.int64, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(_lib.load().b2s_map2dto3d(_ptr(a), _DT_ANY[a.dtype], _ptr(z), _DT_ANY[z.dtype], H, W, B,
                                             float(cx), float(cy), float(fx), float(fy), float(threshold),
                                             _ptr(out), _ptr(npass), _stream()))
    n = int(npass.item())
    if n != B * H * W:          # numpy's message for the reference's np.reshape(u3d, pt2D.shape)
        raise ValueError("cannot reshape array of size %d into shape %s" % (n, str(tuple(a.shape[1:] if squeeze else a.shape))))
    return _from_device(out[0] if squeeze else out, kind)


def remap(src, map1, map2, interpolation=INTER_LINEAR, dst=None, borderMode=0, borderValue=0):
    """cv2.remap(src, mapx, mapy, cv2.INTER_LINEAR), BORDER_CONSTANT 0 (rectified_img_cal.py:253-254;
    SURVEY.md A.7).  src (H,W) / (H,W,3) uint8, or batched (B,H,W) is NOT inferred: pass
    batched colour as (B,H,W,3) and batched gray as (B,H,W,1)."""
    if interpolation != INTER_LINEAR:
        raise error("remap: only INTER_LINEAR is supported", _lib.E_UNSUPPORTED)
    if borderMode != 0 or (np.asarray(borderValue) != 0).any():
        raise error("remap: only BORDER_CONSTANT with value 0 is supported", _lib.E_UNSUPPORTED)
    s, kind = _to_device(src, "src")
    mx, _ = _to_device(map1, "map1")
    my, _ = _to_device(map2, "map2")
    if s.dtype != torch.uint8:
        raise error("remap: only uint8 images are supported", _lib.E_UNSUPPORTED)
    if mx.dtype != torch.float32 or my.dtype != torch.float32 or mx.shape != my.shape or mx.dim() != 2:
        raise error("remap: map1/map2 must be CV_32FC1 maps of the same size")
    if s.dim() == 2:
        B, sH, sW, cn, oshape = 1, s.shape[0], s.shape[1], 1, lambda H, W: (H, W)
    elif s.dim() == 3:
        B, sH, sW, cn, oshape = 1, s.shape[0], s.shape[1], s.shape[2], lambda H, W: (H, W, s.shape[2])
    elif s.dim() == 4:
        B, sH, sW, cn, oshape = s.shape[0], s.shape[1], s.shape[2], s.shape[3], lambda H, W: (s.shape[0], H, W, s.shape[3])
    else:
        raise error("remap: bad source shape")
    H, W = mx.shape
    out = torch.empty(oshape(H, W), dtype=torch.uint8, device=s.device)
    with torch.cuda.device(s.device):
        _lib.check(_lib.load().b2s_remap_u8(_ptr(s), sH, sW, cn, _ptr(mx), _ptr(my), H, W, B, _ptr(out), _stream()))
    res = _from_device(out, kind)
    if dst is not None:
        if isinstance(dst, np.ndarray):
            dst[...] = res
        else:
            dst.copy_(out)
        return dst
    return res


def cvtColor(src, code):
    """cv2.cvtColor(src, cv2.COLOR_BGR2GRAY) for uint8 (disparity_calculation.py:286-287)."""
    if code != COLOR_BGR2GRAY:
        raise error("cvtColor: only COLOR_BGR2GRAY is supported", _lib.E_UNSUPPORTED)
    s, kind = _to_device(src, "src")
    if s.dtype != torch.uint8 or s.dim() < 3 or s.shape[-1] != 3:
        raise error("cvtColor: expected (..., 3) uint8 BGR")
    out = torch.empty(s.shape[:-1], dtype=torch.uint8, device=s.device)
    with torch.cuda.device(s.device):
        _lib.check(_lib.load().b2s_bgr2gray(_ptr(s), out.numel(), _ptr(out), _stream()))
    return _from_device(out, kind)


def medianBlur(src, ksize):
    """cv2.medianBlur(int16 image, 3), replicate border (the blur StereoSGBM.compute applies, A.4)."""
    if ksize != 3:
        raise error("medianBlur: only ksize=3 is supported", _lib.E_UNSUPPORTED)
    s, kind = _to_device(src, "src")
    if s.dtype != torch.int16:
        raise error("medianBlur: only int16 images are supported", _lib.E_UNSUPPORTED)
    squeeze = s.dim() == 2
    if squeeze:
        s = s[None]
    B, H, W = s.shape
    out = torch.empty_like(s)
    with torch.cuda.device(s.device):
        _lib.check(_lib.load().b2s_median3_i16(_ptr(s), H, W, B, _ptr(out), _stream()))
    return _from_device(out[0] if squeeze else out, kind)


def filterSpeckles(img, newVal, maxSpeckleSize, maxDiff, buf=None):
    """cv2.filterSpeckles: in place for torch CUDA tensors; numpy arrays are updated in place too
    and returned (cv2 returns (img, buf))."""
    s, kind = _to_device(img, "img")
    src8 = s if s.dtype == torch.uint8 else None
    if src8 is not None:
        # CV_8UC1 (cv2 accepts it too): the same labelling on an int16 copy; newVal goes through cv2's (uchar) cast
        s = src8.to(torch.int16)
        newVal = int(newVal) & 0xff
    if s.dtype != torch.int16:
        raise error("filterSpeckles: img.type() == CV_8UC1 || img.type() == CV_16SC1", _lib.E_UNSUPPORTED)
    view = s[None] if s.dim() == 2 else s
    B, H, W = view.shape
    labels = torch.empty(B * H * W, dtype=torch.int32, device=s.device)
    counts = torch.empty(B * H * W, dtype=torch.int32, device=s.device)
    with torch.cuda.device(s.device):
        _lib.check(_lib.load().b2s_filter_speckles(_ptr(view), H, W, B, int(newVal), int(maxSpeckleSize), int(maxDiff),
                                                   _ptr(labels), _ptr(counts), _stream()))
    if src8 is not None:
        s = s.to(torch.uint8)
    if kind == "numpy":
        img[...] = s.cpu().numpy()
        return img, buf
    if kind == "torch_cpu":
        img.copy_(s.cpu())
        return img, buf
    if s.data_ptr() != img.data_ptr():
        img.copy_(s)
    return img, buf


def remap_gray(src_bgr, map1, map2):
    """cv2.cvtColor(cv2.remap(src_bgr, map1, map2, INTER_LINEAR), COLOR_BGR2GRAY) in one kernel
    (disparity_calculation.py:207-208 + :286-287).  src (H,W,3) or (B,H,W,3) uint8 ->
    (H',W') or (B,H',W') uint8, byte-identical to the two cv2 calls."""
    s, kind = _to_device(src_bgr, "src")
    if s.dtype != torch.uint8 or s.dim() not in (3, 4) or s.shape[-1] != 3:
        raise error("remap_gray: src must be (H,W,3) or (B,H,W,3) uint8")
    squeeze = s.dim() == 3
    if squeeze:
        s = s[None]
    mx, _ = _to_device(map1, "map1")
    my, _ = _to_device(map2, "map2")
    if mx.dtype != torch.float32 or my.dtype != torch.float32 or mx.shape != my.shape or mx.dim() != 2:
        raise error("remap_gray: map1/map2 must be float32 (H,W) arrays of the same shape")
    B, sH, sW, _c = s.shape
    H, W = mx.shape
    out = torch.empty((B, H, W), dtype=torch.uint8, device=s.device)
    with torch.cuda.device(s.device):
        _lib.check(_lib.load().b2s_remap_gray_u8(_ptr(s), sH, sW, _ptr(mx), _ptr(my), H, W, B, _ptr(out), _stream()))
    return _from_device(out[0] if squeeze else out, kind)


def pipeline(matcher, left_raw, right_raw, maps_left, maps_right, Q=None, *, disp_scale=1.0 / 16.0,
             handleMissingValues=False):
    """The reference's per-pair chain in ONE C call (b2s_pipeline): remap both raw images (BGR input is
    converted to gray in the same kernel) -> matcher.compute -> reprojectImageTo3D(disp * disp_scale, Q).
    Replaces disparity_calculation.py:207-208, 286-289, 302.

    left_raw/right_raw: (H,W) / (H,W,3) uint8 or batched (B,H,W,1) / (B,H,W,3); maps_*: (mapx, mapy) float32.
    Returns (disp int16, xyz float32 or None when Q is None, rect_left, rect_right)."""
    from .engine import StereoSGBM
    if not isinstance(matcher, StereoSGBM):
        raise error("pipeline: matcher must be a StereoSGBM")
    l, kind = _to_device(left_raw, "left")
    r, _ = _to_device(right_raw, "right")
    if l.shape != r.shape or l.dtype != torch.uint8 or r.dtype != torch.uint8:
        raise error("pipeline: left.size() == right.size() && left.type() == right.type() && left.depth() == CV_8U")
    squeeze = l.dim() == 2 or (l.dim() == 3 and l.shape[-1] == 3)
    if l.dim() == 2:
        l, r = l[None, :, :, None], r[None, :, :, None]
    elif l.dim() == 3:
        l, r = l[None], r[None]
    if l.dim() != 4 or l.shape[-1] not in (1, 3):
        raise error("pipeline: images must be (H,W), (H,W,3), (B,H,W,1) or (B,H,W,3)")
    B, sH, sW, cn = l.shape
    mxl, _ = _to_device(maps_left[0], "mapx_l"); myl, _ = _to_device(maps_left[1], "mapy_l")
    mxr, _ = _to_device(maps_right[0], "mapx_r"); myr, _ = _to_device(maps_right[1], "mapy_r")
    for m in (mxl, myl, mxr, myr):
        if m.dtype != torch.float32 or m.shape != mxl.shape or m.dim() != 2:
            raise error("pipeline: the four maps must be float32 (H,W) arrays of one shape")
    H, W = mxl.shape
    dev = l.device
    rl = torch.empty((B, H, W), dtype=torch.uint8, device=dev)
    rr = torch.empty_like(rl)
    disp = torch.empty((B, H, W), dtype=torch.int16, device=dev)
    xyz = torch.empty((B, H, W, 3), dtype=torch.float32, device=dev) if Q is not None else None
    scratch = torch.empty(B, dtype=torch.float32, device=dev) if handleMissingValues else None
    Qp = None
    if Q is not None:
        Qh = np.ascontiguousarray(Q.detach().cpu().numpy() if isinstance(Q, torch.Tensor) else Q, dtype=np.float64)
        if Qh.shape != (4, 4):
            raise error("pipeline: Q.size() == Size(4,4)")
        Qp = Qh.ctypes.data_as(C.POINTER(C.c_double))
    matcher._push()
    with torch.cuda.device(dev):
        _lib.check(_lib.load().b2s_pipeline(
            matcher._h, _ptr(l), _ptr(r), sH, sW, cn, _ptr(mxl), _ptr(myl), _ptr(mxr), _ptr(myr), H, W, B,
            _ptr(rl), _ptr(rr), _ptr(disp), Qp, C.c_float(disp_scale), int(bool(handleMissingValues)),
            _ptr(scratch) if scratch is not None else None, _ptr(xyz) if xyz is not None else None, _stream()))
    outs = [disp, xyz, rl, rr]
    if squeeze:
        outs = [o[0] if o is not None else None for o in outs]
    return tuple(_from_device(o, kind) if o is not None else None for o in outs)
