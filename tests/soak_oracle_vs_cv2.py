"""Randomised soak of the CPU restatement (oracle/) against live cv2, far beyond the committed vectors: every mode,
tiny and saturating penalties, uniquenessRatio up to 99, fully negative ranges, colour, blockSize 1-11.  Not collected by
pytest (a script: it takes minutes); tests/test_oracle_vs_cv2.py holds the cases it found as fixed regressions.

    python tests/soak_oracle_vs_cv2.py [n_cases] [seed0] [workers]
"""
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def draw(k):
    from stereo_match_b200.synth import synth_pair
    from vectors import color_pair
    rng = np.random.default_rng(777000 + k)
    D = int(rng.choice([16, 32, 48, 64, 96, 128, 128, 256]))
    bs = int(rng.choice([1, 1, 3, 5, 7, 9, 11]))
    cn = 3 if rng.random() < 0.15 and bs <= 5 else 1
    cap = int(rng.choice([1, 15, 31, 63]))
    cmax = bs * bs * (2 * (max(cap, 15) | 1) + 63) * cn
    hi = (32767 - 2 * cmax) // 3
    if hi < 16:
        return None
    regime = int(rng.integers(0, 3))                       # tiny / ordinary / as large as the 16-bit domain allows
    P2 = int(rng.integers(3, 120)) if regime == 0 else int(rng.integers(8, min(hi, 3000) + 1)) if regime == 1 else int(rng.integers(hi // 2, hi + 1))
    P1 = int(rng.integers(1, max(P2, 2)))
    minD = int(rng.choice([0, 0, 0, 7, -9, -D // 2, -D]))
    w1 = int(rng.choice([rng.integers(bs, 30), rng.integers(30, 120)]))
    w = w1 + max(D + minD, 0) - min(minD, 0)
    h = int(rng.choice([rng.integers(1, 12), rng.integers(12, 50)]))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=int(rng.choice([-1, 0, 1, 2, 50, 100000])),
              preFilterCap=cap, uniquenessRatio=int(rng.choice([0, 1, 5, 15, 40, 60, 90, 99])), speckleWindowSize=int(rng.choice([0, 0, 30])),
              speckleRange=int(rng.choice([1, 2])), mode=int(rng.integers(0, 4)))
    maxd = max(D + minD, 8)
    l, r = synth_pair(w, h, maxd, seed=k) if cn == 1 else color_pair(w, h, maxd, k)
    if rng.random() < 0.2:                                  # flat, mismatching halves: every path cost saturates
        l = l.copy(); r = r.copy()
        l[:, w // 2:] = 255 - l[:, w // 2:] // 8
        r[:, w // 2:] = r[:, w // 2:] // 8
    return kw, l, r


def one(k):
    import cv2
    from oracle import port
    c = draw(k)
    if c is None:
        return None
    kw, l, r = c
    try:
        want = cv2.StereoSGBM_create(**kw).compute(l, r)
    except cv2.error:
        return None
    got = port.sgbm_compute(l, r, **kw)
    n = int((got != want).sum())
    return (k, n, kw, l.shape) if n else None


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    s0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    nw = int(sys.argv[3]) if len(sys.argv) > 3 else max(1, (os.cpu_count() or 2) // 2)
    with mp.get_context("fork").Pool(nw) as pool:
        res = [x for x in pool.imap_unordered(one, range(s0, s0 + n), chunksize=4) if x]
    for k, npx, kw, shape in sorted(res):
        print("FAIL case %d: %d pixels differ %r %r" % (k, npx, kw, shape))
    print("oracle soak: %d cases, %d failures" % (n, len(res)))
    sys.exit(1 if res else 0)
