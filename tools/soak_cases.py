"""The random case generator of tools/soak_parity.py (shared with the local oracle-vs-cv2 check)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from stereo_match_b200.synth import synth_pair   # noqa: E402
from vectors import color_pair                   # noqa: E402


def gen(k, seed0=0):
    """-> (kw, left, right, cn, batch) or None when the drawn block size leaves no overflow-free P2."""
    rng = np.random.default_rng(31337 + seed0 + k)
    D = int(rng.choice([128, 128, 128, 256]))
    bs = int(rng.choice([1, 3, 5, 5, 7]))
    cn = 3 if rng.random() < 0.15 and bs <= 5 else 1
    cap = int(rng.choice([15, 31, 63]))
    cmax = bs * bs * (2 * (max(cap, 15) | 1) + 63) * cn
    hi = (32767 - 2 * cmax) // 3
    if hi < 16:
        return None
    P2 = int(rng.integers(8, min(hi, 9000)))
    P1 = int(rng.integers(1, P2))
    minD = int(rng.choice([0, 0, 0, 7, -9, -60, -D]))
    w1 = int(rng.choice([rng.integers(bs, 40), rng.integers(40, 130), rng.integers(130, 400)]))
    w = w1 + max(D + minD, 0) - min(minD, 0)
    h = int(rng.choice([rng.integers(1, 20), rng.integers(20, 70), rng.integers(70, 140)]))
    mode = int(rng.choice([0, 1, 1, 3, 2]))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=int(rng.choice([-1, 1, 2, 50])),
              preFilterCap=cap, uniquenessRatio=int(rng.choice([0, 5, 15, 60, 99])), speckleWindowSize=int(rng.choice([0, 60])),
              speckleRange=int(rng.choice([1, 2])), mode=mode)
    maxd = max(D + minD, 8)
    l, r = synth_pair(w, h, maxd, seed=k) if cn == 1 else color_pair(w, h, maxd, k)
    B = int(rng.choice([1, 1, 3]))
    return kw, l, r, cn, B
