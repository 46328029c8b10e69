"""Point cloud behind reprojectImageTo3D: the `points_3d[mask]` / `colors[mask]` step of the reference
(disparity_calculation.py:309-321, try_try.py:91-94) as an ordered device compaction, and the PLY writer of
io_functions.py:28-44 (host file I/O, same bytes)."""
import ctypes as C
import math

import numpy as np
import torch

from . import _lib
from ._lib import error
from .engine import _to_device, _from_device, _ptr, _stream

_DT = {torch.uint8: _lib.DT_U8, torch.int16: _lib.DT_I16, torch.int32: _lib.DT_I32, torch.float32: _lib.DT_F32}


def point_cloud(disparity, points_3d, colors_bgr=None, lo=-math.inf, hi=math.inf):
    """Vertices of the pixels with lo < disparity < hi, in row-major order.

    disparity: (H, W) or (B, H, W), uint8 / int16 / int32 / float32; points_3d: (..., 3) float32 from
    reprojectImageTo3D; colors_bgr: optional (..., 3) uint8 image (the rectified left image) -- returned as RGB,
    like the reference's cvtColor(BGR2RGB) before masking.  NaN / inf coordinates become 0.

    Returns (verts [n, 3] float32, colors [n, 3] uint8 or None) for one image; for a batch a third value
    `offsets` (B + 1 int64): image b owns verts[offsets[b]:offsets[b + 1]].  Same container kind as `disparity`.
    """
    d, kind = _to_device(disparity, "disparity")
    p, _ = _to_device(points_3d, "points_3d")
    if d.dtype not in _DT:
        raise error("point_cloud: disparity must be uint8, int16, int32 or float32", _lib.E_UNSUPPORTED)
    squeeze = d.dim() == 2
    if squeeze:
        d = d[None]
    if d.dim() != 3:
        raise error("point_cloud: disparity must be (H, W) or (B, H, W)")
    B, H, W = d.shape
    if p.dtype != torch.float32 or p.numel() != B * H * W * 3:
        raise error("point_cloud: points_3d must be float32 of shape disparity.shape + (3,)")
    c = None
    if colors_bgr is not None:
        c, _ = _to_device(colors_bgr, "colors_bgr")
        if c.dtype != torch.uint8 or c.numel() != B * H * W * 3:
            raise error("point_cloud: colors_bgr must be uint8 of shape disparity.shape + (3,)")
    lib = _lib.load()
    with torch.cuda.device(d.device):
        verts = torch.empty((B * H * W, 3), dtype=torch.float32, device=d.device)
        rgb = torch.empty((B * H * W, 3), dtype=torch.uint8, device=d.device) if c is not None else None
        offsets = torch.empty(B + 1, dtype=torch.int64, device=d.device)
        scratch = torch.empty(int(lib.b2s_compact_points_scratch_bytes(H, W, B)), dtype=torch.uint8, device=d.device)
        _lib.check(lib.b2s_compact_points(_ptr(p), _ptr(d), _DT[d.dtype], float(lo), float(hi),
                                          _ptr(c) if c is not None else None, H, W, B, _ptr(verts),
                                          _ptr(rgb) if rgb is not None else None, _ptr(offsets), _ptr(scratch), _stream()))
        n = int(offsets[B].item())                      # the one synchronisation: the vertex count
    verts = verts[:n]
    rgb = rgb[:n] if rgb is not None else None
    out = (_from_device(verts, kind), _from_device(rgb, kind) if rgb is not None else None)
    return out if squeeze else out + (_from_device(offsets, kind),)


_PLY_HEADER = ("ply\nformat ascii 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
               "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n")


def write_ply(fn, verts, colors):
    """Same file as the reference's io_functions.write_ply(fn, verts, colors): ASCII PLY, one
    '%f %f %f %d %d %d ' line per vertex."""
    if isinstance(verts, torch.Tensor):
        verts = verts.cpu().numpy()
    if isinstance(colors, torch.Tensor):
        colors = colors.cpu().numpy()
    verts = np.asarray(verts).reshape(-1, 3)
    colors = np.asarray(colors).reshape(-1, 3)
    rows = np.hstack([verts, colors])
    with open(fn, "wb") as f:
        f.write((_PLY_HEADER % len(rows)).encode("utf-8"))
        np.savetxt(f, rows, fmt="%f %f %f %d %d %d ")
