"""Determinism stress of the engine: the 80 random StereoSGBM cases of tests/random_cases.py and the soak regression vectors,
each answered once by cv2 and then computed N times in ONE long process, interleaved with each other and with batches of
other work, fresh handles and reused ones.  A race in a kernel shows up here as a rare mismatch, a stale
shared-memory read in the cost kernels as an overflow flag that comes and goes.
usage: python tools/stress_determinism.py [repetitions]      (GPU box)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cv2                                               # noqa: E402
import stereo_match_b200 as sm                           # noqa: E402
from random_cases import random_case                     # noqa: E402
from vectors import SGBM_SOAK_VECTORS, soak_pair, sgbm_kwargs   # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
cases = []
for seed in range(80):
    kw, l, r = random_case(seed)
    cases.append(("random %d" % seed, kw, l, r))
for i, v in enumerate(SGBM_SOAK_VECTORS):
    l, r = soak_pair(v)
    cases.append(("soak %d" % i, sgbm_kwargs(v[2:]), l, r))
want = [cv2.StereoSGBM_create(**kw).compute(l, r) for _, kw, l, r in cases]
handles = [sm.StereoSGBM_create(**kw) for _, kw, _, _ in cases]
over = []
for (name, kw, l, r), h in zip(cases, handles):          # the overflow flag of a first, quiet run: the reference for the rest
    h.compute(l, r)
    over.append(h.overflowed())
print("cases outside the guaranteed domain (flag set):", [c[0] for c, o in zip(cases, over) if o])
junk = torch.randint(0, 255, (16, 120, 400), dtype=torch.uint8, device="cuda")
big = sm.StereoSGBM_create(numDisparities=128, blockSize=5, P1=600, P2=2400, mode=1)
rng = np.random.default_rng(0)
bad = runs = 0
for rep in range(reps):
    order = rng.permutation(len(cases))
    for j in order:
        name, kw, l, r = cases[j]
        if rng.random() < 0.2:
            big.compute(junk, junk)                      # other kernels in flight before / around
        m = handles[j] if rng.random() < 0.5 else sm.StereoSGBM_create(**kw)
        got = m.compute(l, r)
        runs += 1
        if m.overflowed() != over[j]:
            bad += 1
            print("FLAG rep %d %s: overflow flag %d, first run said %d, %r shape %r" % (rep, name, m.overflowed(), over[j], kw, l.shape), flush=True)
        if over[j]:
            continue
        if not np.array_equal(got, want[j]):
            bad += 1
            ys, xs = np.nonzero(got != want[j])
            print("MISMATCH rep %d %s: %d pixels, rows %r, %r shape %r" % (rep, name, len(ys), sorted(set(ys.tolist()))[:8], kw, l.shape), flush=True)
print("stress: %d runs, %d mismatches" % (runs, bad))
sys.exit(1 if bad else 0)
