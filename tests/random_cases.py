"""Seeded random StereoSGBM cases inside the guaranteed-parity domain, shared by the oracle sweep
(CPU) and the GPU sweep: every mode, gray and colour input, ragged sizes, negative ranges, all post filters."""
import numpy as np

from stereo_match_b200.synth import synth_pair
from vectors import color_pair


def random_case(seed):
    rng = np.random.default_rng(7000 + seed)
    mode = int(rng.integers(0, 4))
    cn = 3 if rng.random() < 0.3 else 1
    D = int(rng.choice([8, 16, 24, 32, 40, 48, 64, 96, 128, 144]))
    minD = int(rng.choice([0, 0, 0, 3, 16, -5, -24]))
    bs = int(rng.choice([1, 3, 5, 7, 9, 11])) if cn == 1 else int(rng.choice([1, 3, 5]))
    cap = int(rng.choice([0, 15, 31, 63]))
    ftzero = max(cap, 15) | 1
    cmax = bs * bs * (2 * ftzero + 63) * cn
    # P2 + largest block cost must stay inside int16 with room for the path sums (A.0): 2*cmax + 3*P2 <= 32767
    p2max = (32767 - 2 * cmax) // 3
    if p2max < 8:
        bs, cmax = 3, 9 * (2 * ftzero + 63) * cn
        p2max = (32767 - 2 * cmax) // 3
    P2 = int(rng.integers(4, max(5, min(p2max, 6000))))
    P1 = int(rng.integers(1, max(2, P2)))
    w = int(rng.integers(max(D + minD, 0) + bs // 2 + 2, max(D + minD, 0) + 180))
    w = max(w, abs(minD) + bs // 2 + 4)
    h = int(rng.choice([1, 2, 5, 9, 17, 33, 40, 65, 97]))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2,
              disp12MaxDiff=int(rng.choice([-1, 0, 1, 2, 10])), preFilterCap=cap,
              uniquenessRatio=int(rng.choice([-1, 0, 5, 10, 15, 40])),
              speckleWindowSize=int(rng.choice([0, 0, 20, 100])), speckleRange=int(rng.choice([1, 2, 4])), mode=mode)
    maxd = max(D + minD, 8)
    if cn == 1:
        l, r = synth_pair(w, h, maxd, seed=9000 + seed)
    else:
        l, r = color_pair(w, h, maxd, 300 + seed)
    return kw, l, r


def random_bm_case(seed):
    """StereoBM as the reference calls it (minDisparity 0) with random window / range / filter settings."""
    rng = np.random.default_rng(8000 + seed)
    D = 16 * int(rng.integers(1, 11))
    bs = int(rng.choice([5, 7, 9, 11, 13, 15, 17, 19, 21, 23, 27]))
    setters = {}
    if rng.random() < 0.7:
        setters["preFilterCap"] = int(rng.integers(1, 64))
    if rng.random() < 0.7:
        setters["textureThreshold"] = int(rng.choice([0, 5, 10, 50, 300]))
    if rng.random() < 0.7:
        setters["uniquenessRatio"] = int(rng.choice([0, 5, 15, 30]))
    if rng.random() < 0.4:
        setters["speckleWindowSize"] = int(rng.choice([10, 50, 200]))
        setters["speckleRange"] = int(rng.choice([-1, 0, 1, 2, 8, 32]))
    w = int(rng.integers(bs + 2, D + bs + 200))
    h = int(rng.integers(bs + 1, 140))
    if w < D + bs and "speckleWindowSize" in setters:
        # cv2 4.13.0 hands back uninitialised memory when no column is valid and a speckle window is set: not a contract
        w += D
    l, r = synth_pair(w, h, max(D, 8), seed=9500 + seed)
    return D, bs, setters, l, r
