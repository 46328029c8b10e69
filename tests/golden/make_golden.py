"""Generates tests/golden/*.npz from live cv2 4.13.0 (the reference's real arithmetic for the hot
path; the reference itself holds no golden vectors).  Inputs are stored
next to the outputs so the fixtures do not depend on cv2 being importable when they are used.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import cv2

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from vectors import (SGBM_VECTORS, SGBM_COLOR_VECTORS, SGBM_SOAK_VECTORS, BM_VECTORS, sgbm_kwargs, sgbm_pair, color_pair,   # noqa: E402
                     soak_pair)
from stereo_match_b200.synth import synth_pair, KITTI_Q                # noqa: E402

assert cv2.__version__ == "4.13.0", cv2.__version__

out = {}
for i in (0, 1, 3, 5, 6, 7, 11, 13, 26, 27, 30, 36, 37, 40):   # 26.. : MODE_HH4, 36.. : MODE_SGBM_3WAY
    v = SGBM_VECTORS[i]
    l, r = sgbm_pair(i, v)
    out["sgbm%d_l" % i], out["sgbm%d_r" % i] = l, r
    out["sgbm%d_v" % i] = np.array(v, np.int32)
    out["sgbm%d_d" % i] = cv2.StereoSGBM_create(**sgbm_kwargs(v)).compute(l, r)
for i, v in enumerate(SGBM_SOAK_VECTORS):                   # the cv2 rules the randomised soaks found (ties, 16-bit wrap, saturation)
    l, r = soak_pair(v)
    out["soak%d_l" % i], out["soak%d_r" % i] = l, r
    out["soak%d_v" % i] = np.array(v[1:], np.int32)
    out["soak%d_d" % i] = cv2.StereoSGBM_create(**sgbm_kwargs(v[2:])).compute(l, r)
for i, seed in SGBM_COLOR_VECTORS[:4]:                      # 3-channel input
    v = SGBM_VECTORS[i]
    l, r = color_pair(v[0], v[1], max(v[2] + v[3], 8), seed)
    out["color%d_l" % i], out["color%d_r" % i] = l, r
    out["color%d_d" % i] = cv2.StereoSGBM_create(**sgbm_kwargs(v)).compute(l, r)
for i in (0, 1, 3):
    w, h, D, bs, setters = BM_VECTORS[i]
    l, r = synth_pair(w, h, D, seed=200 + i)
    m = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
    for k, val in setters.items():
        getattr(m, "set" + k[0].upper() + k[1:])(val)
    out["bm%d_l" % i], out["bm%d_r" % i] = l, r
    out["bm%d_p" % i] = np.array([D, bs, setters.get("preFilterCap", 31), setters.get("textureThreshold", 10),
                                  setters.get("uniquenessRatio", 15)], np.int32)
    out["bm%d_d" % i] = m.compute(l, r)

# StereoBM beyond the defaults (round 2): speckle filter (range unscaled, negative range = off) and minDisparity != 0
# (the right-view matcher of cv2.ximgproc.createRightMatcher).  Parameters as a fixed-order int array.
BMX_KEYS = ("preFilterCap", "textureThreshold", "uniquenessRatio", "speckleWindowSize", "speckleRange", "minDisparity", "disp12MaxDiff")
BMX_DEFAULTS = dict(preFilterCap=31, textureThreshold=10, uniquenessRatio=15, speckleWindowSize=0, speckleRange=0, minDisparity=0,
                    disp12MaxDiff=-1)
for i in range(11, len(BM_VECTORS)):
    w, h, D, bs, setters = BM_VECTORS[i]
    l, r = synth_pair(w, h, D, seed=200 + i)
    m = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
    for k, val in setters.items():
        getattr(m, "set" + k[0].upper() + k[1:])(val)
    out["bmx%d_l" % i], out["bmx%d_r" % i] = l, r
    out["bmx%d_p" % i] = np.array([D, bs] + [setters.get(k, BMX_DEFAULTS[k]) for k in BMX_KEYS], np.int32)
    out["bmx%d_d" % i] = m.compute(l, r)

rng = np.random.default_rng(42)
d16 = rng.integers(-16, 1024, (24, 40)).astype(np.int16)
d16[0, :3] = 0
out["rp_d"] = d16
out["rp_Q"] = KITTI_Q
out["rp_xyz"] = cv2.reprojectImageTo3D(d16, KITTI_Q)
out["rp_xyz_missing"] = cv2.reprojectImageTo3D(d16, KITTI_Q, handleMissingValues=True)
out["rp_xyz_f32"] = cv2.reprojectImageTo3D(d16.astype(np.float32) / 16.0, KITTI_Q.astype(np.float32))

img = rng.integers(0, 256, (48, 64, 3), dtype=np.uint8)
mx = (rng.random((40, 56), dtype=np.float32) * 70 - 3).astype(np.float32)
my = (rng.random((40, 56), dtype=np.float32) * 54 - 3).astype(np.float32)
out["rm_src"], out["rm_mx"], out["rm_my"] = img, mx, my
out["rm_dst3"] = cv2.remap(img, mx, my, cv2.INTER_LINEAR)
out["rm_dst1"] = cv2.remap(np.ascontiguousarray(img[..., 1]), mx, my, cv2.INTER_LINEAR)
out["gray"] = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)

sp = (rng.integers(0, 30, (40, 50)) * 16).astype(np.int16)
sp[rng.random(sp.shape) < 0.3] = -16
out["sp_in"] = sp
for size, diff in ((0, 0), (10, 16), (100, 32)):
    o = sp.copy()
    cv2.filterSpeckles(o, -16, size, diff)
    out["sp_out_%d_%d" % (size, diff)] = o
out["med_in"] = d16
out["med_out"] = cv2.medianBlur(d16, 3)

# PLY bytes from the reference's own writer (io_functions.py:28-44), imported from /root/reference in this container
import importlib.util   # noqa: E402
import tempfile         # noqa: E402
spec = importlib.util.spec_from_file_location("ref_io_functions", "/root/reference/io_functions.py")
ref_io = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref_io)
pv = (rng.standard_normal((37, 3)) * 100).astype(np.float32)
pv[3, 1] = 0.0
pc = rng.integers(0, 256, (37, 3), dtype=np.uint8)
with tempfile.TemporaryDirectory() as td:
    ref_io.write_ply(os.path.join(td, "a.ply"), pv, pc)
    out["ply_bytes"] = np.frombuffer(open(os.path.join(td, "a.ply"), "rb").read(), np.uint8)
out["ply_verts"], out["ply_colors"] = pv, pc

path = os.path.join(HERE, "cv2_4_13_0.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes,", len(out), "arrays")
