"""Worker of tests/test_gpu_poison.py: runs parity cases with every uninitialised device buffer pre-filled with one byte
pattern -- the library's workspaces through B2S_POISON (csrc/api.cu poison_byte), torch.empty / empty_like through the
patch below -- and prints POISON_OK or the cases that differ.  A kernel that reads memory nobody wrote fails here every
time instead of once in a thousand runs (compute-sanitizer's initcheck is not available on the GPU pool).

    B2S_POISON=165 python tests/poison_worker.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

RANDOM = os.environ["B2S_POISON"] == "random"
PAT = 0 if RANDOM else int(os.environ["B2S_POISON"]) & 0xff
_empty, _empty_like = torch.empty, torch.empty_like


def _fill(t):
    if t.is_cuda and t.numel() and t.is_contiguous():
        b = t.view(torch.uint8)
        if RANDOM:
            b.copy_(torch.randint(0, 256, b.shape, dtype=torch.uint8, device=t.device))
        else:
            b.fill_(PAT)
    return t


torch.empty = lambda *a, **k: _fill(_empty(*a, **k))
torch.empty_like = lambda *a, **k: _fill(_empty_like(*a, **k))

import cv2                                         # noqa: E402
import stereo_match_b200 as sm                     # noqa: E402
from stereo_match_b200.synth import synth_pair, KITTI_Q    # noqa: E402
from oracle import port                            # noqa: E402
from vectors import (SGBM_VECTORS, SGBM_SOAK_VECTORS, SGBM_COLOR_VECTORS, BM_VECTORS, sgbm_kwargs, sgbm_pair, soak_pair,   # noqa: E402
                     color_pair)
from random_cases import random_case, random_bm_case       # noqa: E402

bad = []
for i, v in enumerate(SGBM_VECTORS):
    if v[0] * v[1] > 40000:
        continue
    l, r = sgbm_pair(i, v)
    kw = sgbm_kwargs(v)
    mm = sm.StereoSGBM_create(**kw)
    if not np.array_equal(mm.compute(l, r), cv2.StereoSGBM_create(**kw).compute(l, r)):
        bad.append("sgbm vector %d" % i)
    cmax = max(kw["blockSize"], 1) ** 2 * (2 * (max(kw["preFilterCap"], 15) | 1) + 63)
    if mm.overflowed() and 2 * cmax + 3 * max(kw["P2"], 5) <= 32767:      # the flag cannot be genuine for these parameters
        bad.append("sgbm vector %d raises the overflow flag" % i)
for i, v in enumerate(SGBM_SOAK_VECTORS):
    l, r = soak_pair(v)
    kw = sgbm_kwargs(v[2:])
    if not np.array_equal(sm.StereoSGBM_create(**kw).compute(l, r), cv2.StereoSGBM_create(**kw).compute(l, r)):
        bad.append("soak vector %d" % i)
for i, seed in SGBM_COLOR_VECTORS[:6]:
    v = SGBM_VECTORS[i]
    kw = sgbm_kwargs(v)
    l, r = color_pair(v[0], v[1], max(v[2] + v[3], 8), seed)
    if not np.array_equal(sm.StereoSGBM_create(**kw).compute(l, r), cv2.StereoSGBM_create(**kw).compute(l, r)):
        bad.append("colour vector %d" % i)
for seed in range(60):
    kw, l, r = random_case(seed)
    m = sm.StereoSGBM_create(**kw)
    got = m.compute(l, r)
    if m.overflowed():                             # every case of random_cases.py is inside the guaranteed domain
        bad.append("random case %d raises the overflow flag" % seed)
    if not np.array_equal(got, cv2.StereoSGBM_create(**kw).compute(l, r)):
        bad.append("random case %d" % seed)
    if seed % 4 == 0:                              # a batch: handles grow their workspace, bands of several pairs in flight
        L = torch.from_numpy(np.stack([l, l[::-1].copy(), l])).cuda()
        R = torch.from_numpy(np.stack([r, r[::-1].copy(), r])).cuda()
        out = m.compute(L, R).cpu().numpy()
        if m.overflowed() or not (np.array_equal(out[0], got) and np.array_equal(out[2], got)):
            bad.append("random case %d batch" % seed)
for seed in range(20):
    D, bs, setters, l, r = random_bm_case(seed)
    ref = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
    m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
    for name, val in setters.items():
        getattr(ref, "set" + name[0].upper() + name[1:])(val)
        getattr(m, "set" + name[0].upper() + name[1:])(val)
    if not np.array_equal(m.compute(l, r), ref.compute(l, r)):
        bad.append("bm random %d" % seed)
for i, (w, h, D, bs, setters) in enumerate(BM_VECTORS):
    if setters.get("minDisparity", 0) > 0 or w * h > 50000:
        continue
    l, r = synth_pair(w, h, D, seed=200 + i)
    ref = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
    m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
    for name, val in setters.items():
        getattr(ref, "set" + name[0].upper() + name[1:])(val)
        getattr(m, "set" + name[0].upper() + name[1:])(val)
    if not np.array_equal(m.compute(l, r), ref.compute(l, r)):
        bad.append("bm vector %d" % i)
# the chain, the stage operators, the cloud, the WLS filter
rng = np.random.default_rng(5)
l, r = synth_pair(200, 60, 64, seed=9)
yy, xx = np.mgrid[0:60, 0:200].astype(np.float32)
mx, my = (xx + 0.3).astype(np.float32), (yy - 0.2 + 0.001 * xx).astype(np.float32)
kw = dict(minDisparity=0, numDisparities=64, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, preFilterCap=63, uniquenessRatio=15,
          speckleWindowSize=50, speckleRange=2, mode=1)
m = sm.StereoSGBM_create(**kw)
disp, xyz, rl, rr = sm.pipeline(m, l, r, (mx, my), (mx, my), KITTI_Q, handleMissingValues=True)
wl, wr = cv2.remap(l, mx, my, cv2.INTER_LINEAR), cv2.remap(r, mx, my, cv2.INTER_LINEAR)
wd = cv2.StereoSGBM_create(**kw).compute(wl, wr)
wx = cv2.reprojectImageTo3D(wd.astype(np.float32) / 16, KITTI_Q, handleMissingValues=True)
if not (np.array_equal(disp, wd) and np.array_equal(rl, wl) and np.array_equal(xyz.view(np.uint32), wx.view(np.uint32))):
    bad.append("pipeline")
v, c = sm.point_cloud(wd, wx, np.stack([wl] * 3, -1), 0, 600)
mask = (wd > 0) & (wd < 600)
if not (np.array_equal(v.view(np.uint32), np.nan_to_num(wx[mask], nan=0, posinf=0, neginf=0).view(np.uint32)) and len(c) == mask.sum()):
    bad.append("point_cloud")
a, b = wd.copy(), wd.copy()
sm.filterSpeckles(a, -16, 40, 16)
cv2.filterSpeckles(b, -16, 40, 16)
if not np.array_equal(a, b):
    bad.append("filterSpeckles")
if not np.array_equal(sm.medianBlur(wd, 3), cv2.medianBlur(wd, 3)):
    bad.append("medianBlur")
mr = sm.createRightMatcher(m)
f = sm.createDisparityWLSFilter(m)
dl, dr = m.compute(l, r), mr.compute(r, l)
got = f.filter(dl, l, None, dr)
want, _ = port.wls_filter(dl, dr, l, f.getROI(), lambda_=8000.0, sigma=1.0, radius=f.getDepthDiscontinuityRadius())
if not np.array_equal(got, want):
    bad.append("wls")
tag = "random" if RANDOM else str(PAT)
print("POISON_OK pattern %s" % tag if not bad else "POISON_DIFFERS pattern %s: %s" % (tag, ", ".join(bad)))
