"""The random case generator of tools/soak_parity.py (shared with the local oracle-vs-cv2 check)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from stereo_match_b200.synth import synth_pair   # noqa: E402
from vectors import color_pair                   # noqa: E402


def _scene_range(rng, maxD):
    """Disparity scale of the synthetic scene relative to the matcher's range: inside it, reaching its top and beyond
    (the winners then sit in the last disparities, where cv2's SIMD tails live), or only its bottom."""
    return max(int(max(maxD, 8) * float(rng.choice([1.0, 1.0, 1.6, 2.2, 0.4]))), 8)


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
    maxd = _scene_range(rng, D + minD)
    l, r = synth_pair(w, h, maxd, seed=k) if cn == 1 else color_pair(w, h, maxd, k)
    B = int(rng.choice([1, 1, 3]))
    return kw, l, r, cn, B


def gen_wide(k):
    """Every StereoSGBM regime: any numDisparities / blockSize, tiny, ordinary and as-large-as-the-domain-allows penalties,
    ratios up to 99, flat mismatching halves (saturating costs).  -> (kw, left, right) or None."""
    rng = np.random.default_rng(777000 + k)
    D = int(rng.choice([16, 32, 48, 64, 96, 128, 128, 256]))
    bs = int(rng.choice([1, 1, 3, 5, 7, 9, 11]))
    cn = 3 if rng.random() < 0.15 and bs <= 5 else 1
    cap = int(rng.choice([1, 15, 31, 63]))
    cmax = bs * bs * (2 * (max(cap, 15) | 1) + 63) * cn
    hi = (32767 - 2 * cmax) // 3
    if hi < 16:
        return None
    regime = int(rng.integers(0, 3))
    P2 = int(rng.integers(3, 120)) if regime == 0 else int(rng.integers(8, min(hi, 3000) + 1)) if regime == 1 else int(rng.integers(hi // 2, hi + 1))
    P1 = int(rng.integers(1, max(P2, 2)))
    minD = int(rng.choice([0, 0, 0, 7, -9, -D // 2, -D]))
    w1 = int(rng.choice([rng.integers(bs, 30), rng.integers(30, 120)]))
    w = w1 + max(D + minD, 0) - min(minD, 0)
    h = int(rng.choice([rng.integers(1, 12), rng.integers(12, 50)]))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=int(rng.choice([-1, 0, 1, 2, 50, 100000])),
              preFilterCap=cap, uniquenessRatio=int(rng.choice([0, 1, 5, 15, 40, 60, 90, 99])), speckleWindowSize=int(rng.choice([0, 0, 30])),
              speckleRange=int(rng.choice([1, 2])), mode=int(rng.integers(0, 4)))
    maxd = _scene_range(rng, D + minD)
    l, r = synth_pair(w, h, maxd, seed=k) if cn == 1 else color_pair(w, h, maxd, k)
    if rng.random() < 0.2:
        l, r = l.copy(), r.copy()
        l[:, w // 2:] = 255 - l[:, w // 2:] // 8
        r[:, w // 2:] = r[:, w // 2:] // 8
    return kw, l, r


def gen_big(k):
    """numDisparities 272 .. 1024 (32-lane kernels, the per-direction kernels above 512), blockSize up to 15 (generic cost
    kernel), any numDisparities that is not a multiple of 16; few rows to keep cv2's time small.  -> (kw, left, right) or None"""
    rng = np.random.default_rng(888000 + k)
    D = int(rng.choice([264, 272, 300, 320, 400, 512, 520, 528, 640, 1000, 1024, 100, 136, 200]))
    bs = int(rng.choice([1, 3, 5, 7, 9, 13, 15]))
    cn = 3 if rng.random() < 0.2 and bs <= 7 else 1
    cap = int(rng.choice([15, 31, 63]))
    cmax = bs * bs * (2 * (max(cap, 15) | 1) + 63) * cn
    hi = (32767 - 2 * cmax) // 3
    if hi < 16:
        return None
    P2 = int(rng.integers(8, min(hi, 6000) + 1))
    P1 = int(rng.integers(1, P2))
    minD = int(rng.choice([0, 0, 11, -9, -D // 2, -D, -D - 20]))
    w1 = int(rng.choice([rng.integers(bs, 24), rng.integers(24, 100)]))
    w = w1 + max(D + minD, 0) - min(minD, 0)
    h = int(rng.integers(1, 20))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=int(rng.choice([-1, 1, 3, 100000])),
              preFilterCap=cap, uniquenessRatio=int(rng.choice([0, 5, 15, 60, 99])), speckleWindowSize=int(rng.choice([0, 0, 30])),
              speckleRange=int(rng.choice([1, 2])), mode=int(rng.integers(0, 4)))
    maxd = _scene_range(rng, D + minD)
    l, r = synth_pair(w, h, maxd, seed=k) if cn == 1 else color_pair(w, h, maxd, k)
    return kw, l, r


def gen_params(k):
    """cv2's silent parameter fallbacks and odd values: zero / negative penalties, P2 <= P1, negative ratio, preFilterCap
    0 .. 127, zero / negative / huge disp12MaxDiff, negative speckle settings, even block sizes.  -> (kw, left, right)"""
    rng = np.random.default_rng(999000 + k)
    D = int(rng.choice([5, 8, 12, 16, 20, 24, 30, 40, 52, 64, 90, 128]))
    bs = int(rng.choice([-3, 0, 1, 2, 3, 4, 5, 6, 7]))
    cap = int(rng.choice([-5, 0, 1, 14, 15, 16, 30, 63, 64, 100, 127]))
    P1 = int(rng.choice([-7, 0, 1, 2, 50, 400]))
    P2 = int(rng.choice([-7, 0, 1, 2, 3, 50, 51, 400, 2000]))
    minD = int(rng.choice([0, 0, 5, -3, -D]))
    # cv2 only tests W - maxD > blockSize / 2; with a negative minDisparity a computed width W + minD - maxD <= blockSize / 2
    # passes that test and gives NON-DETERMINISTIC output (uninitialised memory; probed): kept out of the domain
    w1 = int(rng.integers((bs if bs > 0 else 5) // 2 + 1 + max(abs(bs), 1), 90))
    w = w1 + max(D + minD, 0) - min(minD, 0)
    h = int(rng.integers(1, 40))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=int(rng.choice([-100, -1, 0, 1, 7, 2 ** 30])),
              preFilterCap=cap, uniquenessRatio=int(rng.choice([-50, -1, 0, 10, 33])), speckleWindowSize=int(rng.choice([-5, 0, 1, 25])),
              speckleRange=int(rng.choice([-2, 0, 1, 3])), mode=int(rng.integers(0, 4)))
    l, r = synth_pair(w, h, _scene_range(rng, D + minD), seed=k)
    return kw, l, r


def gen_bm(k):
    """StereoBM: any window 5..51, numDisparities 16..512, minDisparity <= 0 (what createRightMatcher sets; cv2 itself leaks
    stray pixels for positive ones), every filter.  -> (numDisparities, blockSize, setters, left, right)."""
    rng = np.random.default_rng(555000 + k)
    D = 16 * int(rng.choice([1, 2, 3, 4, 6, 8, 10, 16, 20, 32]))
    bs = int(rng.choice([5, 7, 9, 11, 15, 17, 21, 23, 31, 41, 51]))
    minD = int(rng.choice([0, 0, 0, -1, -7, -D + 1, -D, -D - 9]))
    setters = dict(preFilterCap=int(rng.integers(1, 64)), textureThreshold=int(rng.choice([0, 1, 10, 100, 1000, 5000])),
                   uniquenessRatio=int(rng.choice([0, 1, 5, 15, 50, 90])), minDisparity=minD)
    if rng.random() < 0.4:
        setters["speckleWindowSize"] = int(rng.choice([5, 40, 300]))
        setters["speckleRange"] = int(rng.choice([-1, 0, 1, 4, 40]))
    w = max(D + minD, 0) - min(minD, 0) + bs + int(rng.choice([rng.integers(1, 20), rng.integers(20, 160)]))
    h = int(rng.choice([rng.integers(bs + 1, bs + 8), rng.integers(bs + 8, 120)]))
    l, r = synth_pair(w, h, _scene_range(rng, D + minD), seed=12000 + k)
    return D, bs, setters, l, r


def gen_ops(k):
    """Nasty inputs of the stage operators: map coordinates exactly on pixels / on the 1/32 grid of cv2's fixed-point weights /
    far outside, sparse-to-dense random Q matrices with signed zeros, disparity maps with holes.
    -> dict(src, mx, my, d, size, diff, Q)"""
    rng = np.random.default_rng(424200 + k)
    sh, sw = int(rng.integers(1, 120)), int(rng.integers(1, 200))
    H, W = int(rng.integers(1, 90)), int(rng.integers(1, 160))
    src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
    mx = (rng.random((H, W)) * (sw + 8) - 4).astype(np.float32)
    my = (rng.random((H, W)) * (sh + 8) - 4).astype(np.float32)
    kind = rng.integers(0, 5, (H, W))
    mx = np.where(kind == 1, np.round(mx), mx)
    my = np.where(kind == 1, np.round(my), my)
    mx = np.where(kind == 2, np.round(mx * 32) / 32, mx)
    my = np.where(kind == 2, np.floor(my) + 0.5, my)
    mx = np.where(kind == 3, mx * float(rng.choice([30.0, -30.0, 1e4, 1e6])), mx)
    my = np.where(kind == 3, my * float(rng.choice([1.0, -50.0, 1e6])), my)
    d = (rng.integers(-1, int(rng.choice([3, 60, 2000])), (H, W)) * int(rng.integers(1, 17))).astype(np.int16)
    d[rng.random((H, W)) < float(rng.choice([0.0, 0.3, 0.8]))] = -16
    size, diff = int(rng.choice([0, 1, 7, 50, 400, 100000])), int(rng.choice([0, 1, 16, 32, 1000]))
    Q = rng.standard_normal((4, 4)) * rng.choice([1.0, 100.0, 1e-3], (4, 4)) * (rng.random((4, 4)) < 0.7)    # zeros of both signs
    return dict(src=src, mx=mx.astype(np.float32), my=my.astype(np.float32), d=d, size=size, diff=diff, Q=Q)


def ops_failures(k, ops, cv2):
    """Runs every stage operator of `ops` (the engine: stereo_match_b200; or an oracle adapter with the same names) against
    cv2 on gen_ops(k) -> names of the operators that differ (bitwise; NaN payloads excepted)."""
    c = gen_ops(k)
    src, mx, my, d, Q = c["src"], c["mx"], c["my"], c["d"], c["Q"]
    g = np.ascontiguousarray(src[..., 1])
    fails = []
    if not np.array_equal(ops.remap(src, mx, my, cv2.INTER_LINEAR), cv2.remap(src, mx, my, cv2.INTER_LINEAR)):
        fails.append("remap3")
    if not np.array_equal(ops.remap(g, mx, my, cv2.INTER_LINEAR), cv2.remap(g, mx, my, cv2.INTER_LINEAR)):
        fails.append("remap1")
    if hasattr(ops, "remap_gray") and not np.array_equal(
            ops.remap_gray(src, mx, my), cv2.cvtColor(cv2.remap(src, mx, my, cv2.INTER_LINEAR), cv2.COLOR_BGR2GRAY)):
        fails.append("remap_gray")
    if not np.array_equal(ops.cvtColor(src, cv2.COLOR_BGR2GRAY), cv2.cvtColor(src, cv2.COLOR_BGR2GRAY)):
        fails.append("cvtColor")
    if not np.array_equal(ops.medianBlur(d, 3), cv2.medianBlur(d, 3)):
        fails.append("median")
    a, b = d.copy(), d.copy()
    a = ops.filterSpeckles(a, -16, c["size"], c["diff"])[0]
    cv2.filterSpeckles(b, -16, c["size"], c["diff"])
    if not np.array_equal(a, b):
        fails.append("speckles")
    for Qm in (Q, Q.astype(np.float32)):
        for dd in (d, d.astype(np.float32) / 16, d.astype(np.int32), (d & 0xff).astype(np.uint8)):
            for hm in (False, True):
                want = cv2.reprojectImageTo3D(dd, Qm, handleMissingValues=hm)
                got = ops.reprojectImageTo3D(dd, Qm, handleMissingValues=hm)
                nan = np.isnan(got) & np.isnan(want)
                if not np.array_equal(np.where(nan, 0, got.view(np.uint32)), np.where(nan, 0, want.view(np.uint32))):
                    fails.append("reproject %s Q%s hm%d" % (dd.dtype, Qm.dtype, hm))
    return fails


def gen_chain(k):
    """The reference's per-pair chain (disparity_calculation.py:207-208, 286-289, 302): raw pair (gray or BGR) -> remap through
    two smooth rectification-like maps -> gray -> StereoSGBM -> reprojectImageTo3D.  -> dict or None"""
    rng = np.random.default_rng(616000 + k)
    c = gen_wide(3 * k + 1)
    if c is None:
        return None
    kw, l, r = c
    if l.ndim == 3:                                   # the chain converts to gray before matching
        return None
    H, W = l.shape
    cn = 3 if rng.random() < 0.5 else 1
    sH, sW = H + int(rng.integers(0, 9)), W + int(rng.integers(0, 9))

    def raw(img):
        out = rng.integers(0, 256, (sH, sW), dtype=np.uint8)
        out[:H, :W] = img
        if cn == 1:
            return out
        return np.stack([out, np.roll(out, 1, 1), 255 - out], -1)      # three different channels

    def maps():
        yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
        a, b = rng.normal(0, 2e-3, 2)
        mx = xx + float(rng.normal(0, 1.5)) + a * (yy - H / 2) + 1e-4 * (xx - W / 2) ** 2 / max(W, 1)
        my = yy + float(rng.normal(0, 1.5)) + b * (xx - W / 2)
        return mx.astype(np.float32), my.astype(np.float32)

    f = float(rng.uniform(300, 1500))
    Q = np.array([[1, 0, 0, -rng.uniform(0, W)], [0, 1, 0, -rng.uniform(0, H)], [0, 0, 0, f], [0, 0, 1 / rng.uniform(0.05, 1.0), 0]])
    if rng.random() < 0.3:
        Q[1, 1] = -1.0
        Q[3, 3] = float(rng.normal(0, 0.5))
    return dict(kw=kw, left=raw(l), right=raw(r), maps_left=maps(), maps_right=maps(), Q=Q,
                scale=float(rng.choice([1.0 / 16.0, 1.0])), hm=bool(rng.random() < 0.5), cn=cn)


def structured_pair(rng, w, h, maxd):
    """Image pairs full of exact ties: constant, step edges, checkerboards, stripes, binary noise, identical views, exact
    shifts, half flat / half textured -- the opposite of the textured synthetic scenes."""
    kind = int(rng.integers(0, 9))
    yy, xx = np.mgrid[0:h, 0:w]
    if kind == 0:
        l = np.full((h, w), int(rng.integers(0, 256)), np.uint8)
        r = np.full((h, w), int(rng.integers(0, 256)), np.uint8) if rng.random() < 0.5 else l.copy()
    elif kind == 1:                                        # vertical step edges
        p = int(rng.integers(2, 40))
        l = (((xx // p) % 2) * int(rng.integers(1, 256))).astype(np.uint8)
        r = np.roll(l, -int(rng.integers(0, max(maxd, 1))), 1)
    elif kind == 2:                                        # checkerboard
        p, q = int(rng.integers(1, 12)), int(rng.integers(1, 12))
        l = ((((xx // p) + (yy // q)) % 2) * 255).astype(np.uint8)
        r = np.roll(l, -int(rng.integers(0, max(maxd, 1))), 1)
    elif kind == 3:                                        # horizontal stripes: every disparity looks the same
        l = (((yy // int(rng.integers(1, 6))) % 2) * int(rng.integers(1, 256))).astype(np.uint8)
        r = l.copy()
    elif kind == 4:                                        # binary noise, exact shift
        l = (rng.integers(0, 2, (h, w)) * 255).astype(np.uint8)
        r = np.roll(l, -int(rng.integers(0, max(maxd, 1))), 1)
    elif kind == 5:                                        # identical textured views (disparity 0 everywhere)
        l = rng.integers(0, 256, (h, w), dtype=np.uint8)
        r = l.copy()
    elif kind == 6:                                        # few grey levels: plateaus
        l = (rng.integers(0, 4, (h, w)) * 60).astype(np.uint8)
        r = np.roll(l, -int(rng.integers(0, max(maxd, 1))), 1)
        r[rng.random((h, w)) < 0.05] = 0
    elif kind == 7:                                        # smooth ramp
        l = ((xx * int(rng.integers(1, 5)) + yy) % 256).astype(np.uint8)
        r = np.roll(l, -int(rng.integers(0, max(maxd, 1))), 1)
    else:                                                  # left half flat, right half textured
        l = rng.integers(0, 256, (h, w), dtype=np.uint8)
        l[:, : w // 2] = 128
        r = np.roll(l, -int(rng.integers(0, max(maxd, 1))), 1)
    return np.ascontiguousarray(l), np.ascontiguousarray(r)


def gen_struct(k):
    """gen_wide's parameters on structured_pair images.  -> (kw, left, right) or None"""
    c = gen_wide(k)
    if c is None:
        return None
    kw, l, r = c
    rng = np.random.default_rng(313100 + k)
    h, w = l.shape[:2]
    a, b = structured_pair(rng, w, h, max(kw["numDisparities"] + kw["minDisparity"], 1))
    if l.ndim == 3:
        a, b = np.stack([a, a // 2, 255 - a], -1), np.stack([b, b // 2, 255 - b], -1)
    return kw, a, b


def gen_bm_struct(k):
    D, bs, setters, l, r = gen_bm(k)
    rng = np.random.default_rng(313900 + k)
    a, b = structured_pair(rng, l.shape[1], l.shape[0], max(D + setters["minDisparity"], 1))
    return D, bs, setters, a, b


def gen_large(k):
    """Realistic sizes (up to 1300 x 420) with random parameters and small batches: many bands in flight, the ticket order and
    the progress publication of the band pipeline under different timings.  -> (kw, left, right, cn, batch)"""
    rng = np.random.default_rng(474700 + k)
    D = int(rng.choice([64, 96, 128, 128, 160, 256]))
    bs = int(rng.choice([3, 5, 5, 7]))
    cn = 3 if rng.random() < 0.1 else 1
    cap = int(rng.choice([31, 63]))
    cmax = bs * bs * (2 * (max(cap, 15) | 1) + 63) * cn
    hi = (32767 - 2 * cmax) // 3
    if hi < 100:                                           # colour with a large block leaves no overflow-free P2
        cn, cmax = 1, cmax // 3
        hi = (32767 - 2 * cmax) // 3
    P2 = int(rng.integers(50, min(hi, 5000)))
    P1 = int(rng.integers(1, P2))
    minD = int(rng.choice([0, 0, 0, 16, -20]))
    w = int(rng.integers(max(D + minD, 0) + 200, 1300))
    h = int(rng.integers(60, 420))
    kw = dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=int(rng.choice([-1, 1, 2])),
              preFilterCap=cap, uniquenessRatio=int(rng.choice([0, 5, 15])), speckleWindowSize=int(rng.choice([0, 100])),
              speckleRange=int(rng.choice([1, 2])), mode=int(rng.choice([0, 1, 1, 2, 3])))
    maxd = _scene_range(rng, D + minD)
    l, r = synth_pair(w, h, maxd, seed=k) if cn == 1 else color_pair(w, h, maxd, k)
    return kw, l, r, cn, int(rng.choice([1, 2, 3]))
