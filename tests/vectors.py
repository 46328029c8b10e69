"""Parity vectors shared by the oracle tests and the GPU parity tests."""
from stereo_match_b200.synth import synth_pair

# (w, h, minD, D, blockSize, P1, P2, disp12MaxDiff, preFilterCap, uniquenessRatio, mode, speckleWindow, speckleRange)
SGBM_VECTORS = [
    (96, 64, 0, 16, 5, 600, 2400, 1, 63, 15, 0, 0, 0),
    (96, 64, 0, 16, 5, 600, 2400, 1, 63, 15, 1, 0, 0),
    (120, 70, 0, 32, 3, 72, 288, 1, 63, 10, 0, 0, 0),
    (120, 70, 0, 32, 7, 72, 288, 2, 31, 0, 1, 0, 0),
    (131, 77, 4, 32, 9, 100, 1000, 1, 0, 5, 1, 0, 0),
    (131, 77, -8, 32, 5, 100, 1000, -1, 10, 5, 0, 0, 0),
    (131, 77, -40, 32, 5, 100, 1000, 1, 10, 5, 1, 0, 0),
    (150, 90, 0, 48, 11, 968, 3872, 1, 63, 15, 1, 50, 2),
    (150, 90, 0, 48, 1, 8, 32, 1, 63, 15, 0, 100, 1),
    (150, 90, 0, 16, 11, 50, 200, 20, 1, 0, 0, 0, 0),       # large block, inside the guaranteed domain
    (150, 90, 0, 48, 5, 0, 0, 0, 0, -1, 0, 0, 0),           # every silent fallback
    (150, 90, 0, 48, 5, 3000, 12000, 1, 63, 15, 1, 0, 0),   # int16 pressure
    (150, 90, 0, 24, 5, 100, 400, 1, 63, 15, 0, 0, 0),      # D not a multiple of 16
    (150, 90, 0, 40, 5, 100, 400, 1, 63, 15, 1, 0, 0),
    (200, 60, 0, 160, 5, 600, 2400, 1, 63, 15, 0, 0, 2),    # settings.ini as shipped (D=160)
    (330, 40, 0, 272, 3, 72, 288, 1, 63, 10, 1, 0, 0),      # D > 256: 32-lane kernels
    (70, 40, 0, 64, 5, 600, 2400, 1, 63, 15, 1, 0, 0),      # width1 = 6
    (40, 3, 0, 16, 5, 600, 2400, 1, 63, 15, 1, 0, 0),       # H smaller than the block
    (1, 1, -1, 1, 1, 8, 32, 1, 63, 15, 0, 0, 0),            # one pixel, one disparity
    (640, 24, 0, 528, 3, 72, 288, 1, 63, 10, 1, 0, 0),      # D > 512: per-direction kernels
    (300, 200, 0, 96, 5, 600, 2400, 1, 63, 15, 1, 60, 2),   # 7 bands of 32 rows, R = 6
    (300, 200, 0, 96, 5, 600, 2400, 1, 63, 15, 0, 0, 0),
    (420, 70, 0, 200, 3, 100, 800, 1, 63, 10, 1, 0, 0),     # G = 16, padded (Dp = 256), 5 bands
    (23, 50, 0, 16, 3, 100, 800, 1, 63, 10, 1, 0, 0),       # width1 = 7 < one staging chunk
    (180, 50, 0, 128, 5, 600, 2400, 1, 63, 10, 1, 0, 0),    # R = 8, S summed without per-direction saturation
    (180, 50, 0, 128, 5, 1000, 8000, 1, 63, 10, 1, 0, 0),   # R = 8, large P2: saturating accumulation
    # MODE_HH4 (-> , <- , down, up; cv2 4.13.0 leaves the block cost of the last blockSize/2 rows at zero)
    (150, 90, 0, 48, 5, 100, 400, 1, 63, 10, 3, 0, 0),
    (131, 77, 4, 32, 9, 100, 1000, 1, 0, 5, 3, 50, 2),
    (200, 120, -8, 64, 3, 72, 288, 2, 63, 0, 3, 0, 0),
    (300, 200, 0, 96, 5, 600, 2400, 1, 63, 15, 3, 60, 2),   # 7 bands, R = 6
    (40, 3, 0, 16, 5, 600, 2400, 1, 63, 15, 3, 0, 0),       # H smaller than the block: rows 1.. have zero cost
    (60, 1, 0, 16, 5, 600, 2400, 1, 63, 15, 3, 0, 0),       # one row
    (150, 90, 0, 40, 1, 50, 200, 1, 63, 15, 3, 0, 0),       # blockSize 1: no zeroed rows, padded D
    (180, 50, 0, 128, 5, 600, 2400, 1, 63, 10, 3, 0, 0),    # R = 8 (permuted records)
    (330, 40, 0, 272, 3, 72, 288, 1, 63, 10, 3, 0, 0),      # 32-lane kernels
    (640, 24, 0, 528, 3, 72, 288, 1, 63, 10, 3, 0, 0),      # D > 512: per-direction kernels
    # MODE_SGBM_3WAY (-> , down, <- inside 4 fixed stripes that restart the block sum and the downward path;
    # uniqueness test with <=)
    (300, 200, 0, 64, 5, 100, 400, 1, 63, 10, 2, 0, 0),
    (150, 90, 0, 48, 5, 100, 400, 1, 63, 0, 2, 40, 2),      # uniquenessRatio 0: no uniqueness test at all
    (131, 77, 4, 32, 9, 100, 1000, 1, 0, 5, 2, 0, 0),
    (200, 121, -8, 64, 3, 72, 288, 2, 63, 15, 2, 40, 2),
    (40, 3, 0, 16, 5, 600, 2400, 1, 63, 15, 2, 0, 0),       # stripes of one row: clamped starts shift cv2's rows
    (60, 1, 0, 16, 5, 600, 2400, 1, 63, 15, 2, 0, 0),
    (120, 7, 0, 16, 3, 100, 400, 1, 63, 10, 2, 0, 0),
    (120, 12, 0, 16, 5, 100, 400, 1, 63, 10, 2, 0, 0),
    (100, 25, 0, 16, 9, 100, 400, 1, 63, 10, 2, 0, 0),
    (150, 90, 0, 24, 1, 8, 32, 1, 63, 15, 2, 0, 0),         # equality cases of the uniqueness test
    (150, 90, 0, 48, 5, 0, 0, 0, 0, -1, 2, 0, 0),           # every silent fallback
    (180, 50, 0, 128, 5, 600, 2400, 1, 63, 10, 2, 0, 0),    # R = 8
    (330, 40, 0, 272, 3, 72, 288, 1, 63, 10, 2, 0, 0),      # 32-lane kernels
    (640, 24, 0, 528, 3, 72, 288, 1, 63, 10, 2, 0, 0),      # D > 512: per-direction kernels
    (150, 90, 0, 40, 13, 20, 80, 1, 31, 15, 2, 0, 0),       # blockSize 13: generic cost kernel
]

# 3-channel input (reference: try_try.py:69-81): (index into SGBM_VECTORS whose parameters are used, seed)
SGBM_COLOR_VECTORS = [(0, 1), (1, 2), (3, 3), (4, 4), (5, 5), (8, 6), (12, 7), (22, 8), (24, 9), (26, 10), (27, 11), (36, 12),
                      (37, 13), (40, 14), (47, 15), (46, 16)]


def color_pair(w, h, maxd, seed):
    """Correlated colour channels: one synthetic pair plus per-channel mixtures of two others."""
    import numpy as np
    a, b, c = (synth_pair(w, h, maxd, seed=1000 + 3 * seed + k) for k in range(3))

    def mix(k):
        return [a[k], np.clip(a[k] * 0.8 + b[k] * 0.2, 0, 255).astype(np.uint8),
                np.clip(a[k] * 0.7 + c[k] * 0.3, 0, 255).astype(np.uint8)]
    return np.stack(mix(0), -1), np.stack(mix(1), -1)


# cv2 4.13.0 raises cv2.error for these (width - maxD <= blockSize/2, stereosgbm.cpp:511)
SGBM_ERROR_VECTORS = [
    (50, 30, 0, 64, 5, 600, 2400, 1, 63, 15, 0, 0, 0),
    (66, 30, 0, 64, 5, 600, 2400, 1, 63, 15, 1, 0, 0),
]

# outside the guaranteed-parity domain (cv2's own SIMD and scalar code disagree there,
# SURVEY.md A.0): the engine must raise its overflow flag.  disparity_test.py:165-177
SGBM_UNPINNED_VECTORS = [
    (150, 90, 0, 16, 23, 222, 887, 20, 1, 0, 0, 0, 0),
]


# Found by the randomised soaks (tests/soak_oracle_vs_cv2.py, tools/soak_parity.py) in round 2: three cv2 4.13.0 rules the
# vectors above never reached.  (pair kind, seed) + the 13 fields of SGBM_VECTORS.
#   ties  : MODE_SGBM_3WAY picks a tied minimum on 8 SIMD lanes (last minimum of each lane, smallest position over lanes)
#   wrap  : MODE_SGBM_3WAY compares S against (int16)(100 * minS / (100 - ratio) + 1): wraps for a large ratio / large costs
#   sat   : every S of a pixel saturated at 32767 -> cv2's strict search finds no winner -> INVALID
SGBM_SOAK_VECTORS = [
    ("synth", 12, 159, 1, 0, 128, 1, 37, 47, 100000, 63, 0, 2, 0, 0),        # ties, one row, R = 8
    ("synth", 1, 150, 9, 0, 64, 1, 3, 8, 1, 63, 0, 2, 0, 0),                 # ties, R = 4
    ("synth", 21, 70, 9, 0, 16, 1, 3, 8, 1, 63, 0, 2, 0, 0),                 # ties, R = 2
    ("synth", 41, 140, 9, 0, 96, 1, 3, 8, 1, 15, 0, 2, 0, 0),                # ties, R = 6
    ("synth", 339, 313, 8, 0, 256, 1, 3, 7, 1, 15, 0, 2, 0, 0),              # ties, 16 lanes
    ("synth", 50, 330, 9, 0, 272, 1, 3, 7, 1, 15, 0, 2, 0, 0),               # ties, 32 lanes, padded D
    ("synth", 5, 560, 9, 0, 528, 1, 3, 7, 1, 15, 0, 2, 0, 0),                # ties, D > 512: k_wta
    ("synth", 7, 272, 26, 7, 128, 5, 2000, 5309, 1, 63, 99, 2, 0, 0),        # wrap, R = 8
    ("synth", 8, 150, 30, 0, 48, 5, 1500, 5000, 1, 63, 99, 2, 0, 0),         # wrap, R = 4 padded
    ("synth", 10, 580, 12, 0, 528, 5, 1500, 5000, 1, 63, 99, 2, 0, 0),       # wrap, k_wta
    ("synth", 4, 342, 74, -60, 128, 1, 68, 229, -1, 15, 99, 2, 60, 2),       # wrap and ties, negative range, speckle filter
    ("synth", 535, 157, 98, 0, 128, 7, 3337, 4540, 50, 63, 0, 1, 0, 0),      # sat, MODE_HH, R = 8 (saturating accumulation)
    ("synth", 101, 120, 60, 0, 48, 7, 3337, 4540, 50, 63, 0, 1, 0, 0),       # sat, R = 4
    ("ramp", 3, 100, 40, 0, 48, 9, 10, 100, 50, 63, 0, 1, 0, 0),             # sat: half of the image has no winner at all
    ("ramp", 3, 180, 40, 0, 128, 9, 10, 100, 50, 63, 0, 1, 0, 0),
    ("ramp", 3, 100, 40, 0, 48, 9, 10, 100, 50, 63, 10, 1, 30, 2),
    ("ramp", 3, 180, 40, 0, 128, 9, 10, 100, 50, 63, 0, 2, 0, 0),            # 3WAY on the same pair: it does find winners
    ("ramp", 4, 620, 24, 0, 528, 9, 10, 100, 50, 63, 0, 1, 0, 0),            # sat, per-direction kernels + k_wta
    # MODE_SGBM_3WAY with numDisparities % 8 != 0 (cv2's scalar tail: the in-place paths take a wrong d - 1 neighbour at
    # d = 8 * (D / 8), the tail continues the winner search strictly and uses the 32-bit uniqueness rule); "top" scenes
    # put the true disparities into the last cells of the range, where the tail lives
    ("top", 1, 52, 19, 0, 12, 3, 100, 800, 100000, 31, 0, 2, 0, 0),
    ("top", 2, 60, 19, 0, 20, 3, 100, 800, 1, 31, 15, 2, 0, 0),
    ("top", 3, 70, 19, 0, 30, 3, 8, 32, 1, 31, 15, 2, 0, 0),
    ("top", 4, 45, 19, 0, 5, 3, 8, 32, 1, 31, 10, 2, 0, 0),                  # fewer than 8 disparities: no SIMD part at all
    ("top", 5, 140, 23, 0, 100, 5, 104, 5205, 100000, 63, 0, 2, 0, 0),
    ("top", 6, 184, 15, -50, 100, 7, 20, 2948, -1, 15, 10, 2, 0, 0),
    ("top", 7, 340, 11, 0, 300, 3, 72, 288, 1, 63, 10, 2, 30, 2),
    ("top", 8, 590, 15, 0, 535, 3, 8, 32, 1, 63, 0, 2, 0, 0),               # D > 512 and a tail
    ("top", 9, 140, 23, 0, 100, 5, 104, 5205, 1, 63, 5, 1, 0, 0),            # the other modes have no such tail
    # MODE_SGBM_3WAY: blockSize <= 0 means a 3-pixel window (5 elsewhere), and the stripe overlap takes blockSize as given
    ("synth", 131, 212, 36, 0, 128, -3, 0, 400, -1, -5, -50, 2, 0, 0),
    ("synth", 782, 90, 37, 0, 24, 0, 0, 2, 7, 64, 0, 2, 0, -2),
    ("synth", 440, 108, 54, 0, 40, -3, 400, 2000, 0, 30, 33, 2, 0, 1),
    ("synth", 441, 108, 54, 0, 40, -3, 400, 2000, 0, 30, 33, 1, 0, 1),       # MODE_HH: the 5-pixel fallback
    # MODE_HH4 has no "image too small" test in cv2: it computes whatever width is left, all INVALID when none is
    ("synth", 60, 34, 30, 0, 32, 5, 100, 400, 1, 63, 5, 3, 0, 0),            # width1 = 2 <= blockSize / 2
    ("synth", 61, 33, 30, 0, 32, 9, 100, 400, 1, 63, 5, 3, 20, 2),           # width1 = 1
    ("synth", 62, 20, 12, 0, 32, 5, 100, 400, 1, 63, 5, 3, 0, 0),            # W < numDisparities: nothing computable
    ("synth", 63, 32, 12, 0, 32, 5, 100, 400, 1, 63, 5, 3, 0, 0),            # width1 = 0
]


def soak_pair(v):
    """The image pair of one SGBM_SOAK_VECTORS entry.  "ramp": the lower half is a bright rising saw-tooth against a dark
    falling one, a large cost at EVERY disparity; the upper half an ordinary matching pair."""
    import numpy as np
    kind, seed, w, h, minD, D = v[:6]
    if kind == "synth":
        return synth_pair(w, h, max(D + minD, 8), seed=seed)
    if kind == "top":
        return synth_pair(w, h, max(int(2.5 * (D + minD)), 8), seed=seed)    # the middle rows sit at the top of the range
    rng = np.random.default_rng(9000 + seed)
    x, y = np.arange(w)[None, :], np.arange(h)[:, None]
    l = (192 + ((9 * x + 2 * y) % 64) + rng.integers(0, 3, (h, w)) - 1).clip(0, 255).astype(np.uint8)
    r = (63 - ((9 * x + 3 * y) % 64) + rng.integers(0, 3, (h, w)) + 1).clip(0, 255).astype(np.uint8)
    a, b = synth_pair(w, h, 8, seed=seed)
    l[:h // 2], r[:h // 2] = a[:h // 2], b[:h // 2]
    return l, r


def sgbm_kwargs(v):
    w, h, minD, D, bs, P1, P2, lr, cap, uq, mode, spw, spr = v
    return dict(minDisparity=minD, numDisparities=D, blockSize=bs, P1=P1, P2=P2, disp12MaxDiff=lr,
                preFilterCap=cap, uniquenessRatio=uq, speckleWindowSize=spw, speckleRange=spr, mode=mode)


def sgbm_pair(i, v):
    w, h, minD, D = v[:4]
    return synth_pair(w, h, max(D + minD, 8), seed=100 + i)


# (w, h, D, blockSize, setters)
BM_VECTORS = [
    (100, 60, 16, 5, {}),
    (100, 61, 16, 7, {}),                                   # odd height
    (140, 80, 32, 15, {}),
    (140, 80, 32, 9, dict(preFilterCap=63, textureThreshold=0, uniquenessRatio=0)),
    (90, 50, 64, 21, {}),
    (40, 30, 32, 5, {}),                                    # a handful of valid columns
    (30, 30, 32, 5, {}),                                    # no valid column
    (120, 60, 16, 23, {}),                                  # blockSize > 21: generic kernel
    (300, 250, 48, 11, {}),                                 # several column strips and row bands
    (260, 120, 160, 13, dict(uniquenessRatio=5)),           # 10 disparity warps
    (200, 110, 16, 19, dict(textureThreshold=200)),         # one warp per CTA
    (220, 130, 32, 9, dict(speckleWindowSize=80, speckleRange=2)),      # speckle filter: range NOT scaled by 16
    (220, 130, 32, 9, dict(speckleWindowSize=40, speckleRange=0)),
    (220, 130, 32, 9, dict(speckleWindowSize=80, speckleRange=-1)),     # negative range: filter skipped
    # minDisparity != 0: the right-view matcher of cv2.ximgproc.createRightMatcher is StereoBM over [-(D-1), 0]
    (150, 60, 32, 9, dict(minDisparity=-31, textureThreshold=0, uniquenessRatio=0, disp12MaxDiff=1000000)),
    (150, 60, 48, 15, dict(minDisparity=-47)),
    (150, 60, 16, 7, dict(minDisparity=-40)),                           # range entirely negative (rofs > 0)
    (150, 60, 32, 23, dict(minDisparity=-31)),                          # generic kernel
    (640, 64, 512, 51, {}),                                             # generic kernel, 16-column tiles (shared memory)
    (400, 70, 320, 41, dict(uniquenessRatio=5, textureThreshold=0)),
]

# full-size configurations of BASELINE.json
CFG1 = dict(w=640, h=480, seed=1, kw=dict(minDisparity=0, numDisparities=64, blockSize=5, P1=600, P2=2400,
            disp12MaxDiff=1, uniquenessRatio=15, speckleWindowSize=0, speckleRange=2, preFilterCap=63, mode=0))
CFG2 = dict(w=1280, h=720, seed=2, D=128, bs=15)
CFG3 = dict(w=1242, h=375, seed=1, kw=dict(minDisparity=0, numDisparities=128, blockSize=5, P1=600, P2=2400,
            disp12MaxDiff=1, uniquenessRatio=15, speckleWindowSize=100, speckleRange=2, preFilterCap=63, mode=1))
CFG4 = dict(w=2864, h=1988, seed=4, kw=dict(minDisparity=0, numDisparities=256, blockSize=5, P1=600, P2=2400,
            disp12MaxDiff=1, uniquenessRatio=15, speckleWindowSize=100, speckleRange=2, preFilterCap=63, mode=1))


# reprojectImageTo3D with dense Q matrices holding zeros of BOTH signs and negative first-column entries (found by the
# operator soak): cv2's 4x4 product starts its accumulator at +0.0, which decides the sign of a zero or infinite result
def reproject_q_cases():
    import numpy as np
    rng = np.random.default_rng(8)
    out = []
    for k in range(6):
        Q = rng.standard_normal((4, 4)) * rng.choice([1.0, 100.0, 1e-3], (4, 4)) * (rng.random((4, 4)) < 0.6)
        if k == 0:
            Q = np.array([[42.3, 161.9, 109.6, -0.0], [-44.7, 1.14, 6.84, -0.087], [-0.0, -7.8e-4, -0.0, -0.0], [8.3e-4, 18.6, -0.0, -1.29]])
        if k == 1:
            Q = np.array([[-1.4e-3, -0.97, -0.71, -0.0], [-92.6, 0.33, -8.4e-4, -1.1e-4], [9.19, -0.0, 1.9, 151.4], [-8.2e-4, -2.04, -0.0, -0.0]])
        d = (rng.integers(-1, 60, (23, 41)) * 8).astype(np.int16)
        d[rng.random(d.shape) < 0.3] = -16
        out.append((Q, d))
    return out
