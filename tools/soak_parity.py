"""Randomised parity soak of the engine against live cv2: many more shapes / parameters than the test suite (numDisparities 128 /
256, every mode, ratios up to 99, penalties up to the edge of the 16-bit domain, fully negative ranges, colour, batches).
usage: python tools/soak_parity.py [n_cases] [seed0] [sgbm|wide|big|params|bm|ops|chain]      (GPU box; prints failures, exits 1 on any)
sgbm: the numDisparities 128 / 256 kernels (soak_cases.gen); wide: every regime (gen_wide); big: numDisparities up to 1024,
blocks up to 15 (gen_big); params: cv2's silent fallbacks and odd values (gen_params); bm: StereoBM (gen_bm); ops: remap, cvtColor,
medianBlur, filterSpeckles, reprojectImageTo3D; chain: the fused per-pair chain (sm.pipeline) against the separate cv2 calls.

Round 2: 600 cases found 25 mismatches, all three of them cv2 rules the restatement did not have (tests/vectors.py
SGBM_SOAK_VECTORS); 0 of 600 after."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cv2
import numpy as np
import torch

from soak_cases import gen, gen_wide, gen_big, gen_params, gen_bm, gen_chain, ops_failures

import stereo_match_b200 as sm

n, seed0 = (int(sys.argv[1]) if len(sys.argv) > 1 else 300), (int(sys.argv[2]) if len(sys.argv) > 2 else 0)
which = sys.argv[3] if len(sys.argv) > 3 else "sgbm"
bad = ran = 0


for k in range(seed0, seed0 + n):
    if which == "ops":
        f = ops_failures(k, sm, cv2)
        ran += 1
        if f:
            bad += 1
            print("FAIL ops case %d: %s" % (k, ", ".join(f)))
        continue
    if which == "chain":
        c = gen_chain(k)
        if c is None:
            continue
        try:
            ref = cv2.StereoSGBM_create(**c["kw"])
            rl = cv2.remap(c["left"], c["maps_left"][0], c["maps_left"][1], cv2.INTER_LINEAR)
            rr = cv2.remap(c["right"], c["maps_right"][0], c["maps_right"][1], cv2.INTER_LINEAR)
            if c["cn"] == 3:
                rl, rr = cv2.cvtColor(rl, cv2.COLOR_BGR2GRAY), cv2.cvtColor(rr, cv2.COLOR_BGR2GRAY)
            wd = ref.compute(rl, rr)
        except cv2.error:
            continue
        wx = cv2.reprojectImageTo3D(wd if c["scale"] == 1.0 else wd.astype(np.float32) * np.float32(c["scale"]), c["Q"],
                                    handleMissingValues=c["hm"])
        m = sm.StereoSGBM_create(**c["kw"])
        disp, xyz, grl, grr = sm.pipeline(m, c["left"], c["right"], c["maps_left"], c["maps_right"], c["Q"], disp_scale=c["scale"],
                                          handleMissingValues=c["hm"])
        if m.overflowed():
            continue
        ran += 1
        f = []
        if not (np.array_equal(grl, rl) and np.array_equal(grr, rr)):
            f.append("rectified")
        if not np.array_equal(disp, wd):
            f.append("disparity")
        nan = np.isnan(xyz) & np.isnan(wx)
        if not np.array_equal(np.where(nan, 0, xyz.view(np.uint32)), np.where(nan, 0, wx.view(np.uint32))):
            f.append("xyz")
        if f:
            bad += 1
            print("FAIL chain case %d: %s, %r cn %d scale %g hm %d shape %r" % (k, ", ".join(f), c["kw"], c["cn"], c["scale"], c["hm"], c["left"].shape))
        continue
    if which == "bm":
        D, bs, setters, l, r = gen_bm(k)
        ref = cv2.StereoBM_create(numDisparities=D, blockSize=bs)
        m = sm.StereoBM_create(numDisparities=D, blockSize=bs)
        for name, val in setters.items():
            getattr(ref, "set" + name[0].upper() + name[1:])(val)
            getattr(m, "set" + name[0].upper() + name[1:])(val)
        want = ref.compute(l, r)
        if "speckleWindowSize" in setters and not (want != (setters["minDisparity"] - 1) * 16).any():
            continue              # cv2 returns uninitialised memory when nothing is valid and a speckle window is set
        got = m.compute(l, r)
        ran += 1
        if not np.array_equal(got, want):
            bad += 1
            print("FAIL bm case %d: %d pixels differ, D %d bs %d %r shape %r" % (k, int((got != want).sum()), D, bs, setters, l.shape))
        continue
    c = {"sgbm": gen, "wide": gen_wide, "big": gen_big, "params": gen_params}[which](k)
    if c is None:
        continue
    kw, l, r = c[:3]
    cn, B = (c[3], c[4]) if which == "sgbm" else (l.ndim == 3 and 3 or 1, 1 + 2 * (k % 3 == 0))
    try:
        want = cv2.StereoSGBM_create(**kw).compute(l, r)
    except cv2.error:
        try:
            sm.StereoSGBM_create(**kw).compute(l, r)
        except sm.error:
            continue
        bad += 1
        print("FAIL case %d: cv2 raises, the engine does not, %r shape %r" % (k, kw, l.shape))
        continue
    try:
        m = sm.StereoSGBM_create(**kw)
        m.compute(l, r)
    except sm.error as ex:
        bad += 1
        print("FAIL case %d: the engine raises (%s), cv2 does not, %r shape %r" % (k, str(ex)[:80], kw, l.shape))
        continue
    if B == 1:
        got = m.compute(l, r)
    else:
        L = torch.from_numpy(np.stack([l] * B)).cuda()
        R = torch.from_numpy(np.stack([r] * B)).cuda()
        out = m.compute(L, R).cpu().numpy()
        got = out[B - 1]
        if not all(np.array_equal(out[i], out[0]) for i in range(B)):
            print("BATCH MISMATCH", k, kw, l.shape)
            bad += 1
    if m.overflowed():
        continue
    ran += 1
    if not np.array_equal(got, want):
        bad += 1
        print("FAIL case %d: %d pixels differ, %r shape %r cn %d batch %d" % (k, int((got != want).sum()), kw, l.shape, cn, B))
print("soak %s: %d cases run, %d failures" % (which, ran, bad))
sys.exit(1 if bad else 0)
