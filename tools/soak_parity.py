"""Randomised parity soak of the engine against live cv2: many more shapes / parameters than the test suite (numDisparities 128 /
256, every mode, ratios up to 99, penalties up to the edge of the 16-bit domain, fully negative ranges, colour, batches).
usage: python tools/soak_parity.py [n_cases] [seed0]      (GPU box; prints failures, exits 1 on any)

Round 2: 600 cases found 25 mismatches, all three of them cv2 rules the restatement did not have (tests/vectors.py
SGBM_SOAK_VECTORS); 0 of 600 after."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import cv2
import numpy as np
import torch

from soak_cases import gen

import stereo_match_b200 as sm

n, seed0 = (int(sys.argv[1]) if len(sys.argv) > 1 else 300), (int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for k in range(n):
    c = gen(k, seed0)
    if c is None:
        continue
    kw, l, r, cn, B = c
    try:
        want = cv2.StereoSGBM_create(**kw).compute(l, r)
    except cv2.error:
        continue
    m = sm.StereoSGBM_create(**kw)
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
    if not np.array_equal(got, want):
        bad += 1
        print("FAIL case %d: %d pixels differ, %r shape %r cn %d batch %d" % (k, int((got != want).sum()), kw, l.shape, cn, B))
print("soak: %d cases, %d failures" % (n, bad))
sys.exit(1 if bad else 0)
