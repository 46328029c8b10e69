"""GPU parity of the lean wavefront kernel (stereo_match_b200/csrc/sgbm_wave8.cu): its domain is numDisparities = 128
(8 lanes per pixel) and 256 (16 lanes), MODE_HH (both sweeps) and MODE_SGBM (first sweep), sums that cannot overflow.
Checked bit-exactly against live cv2 4.13.0 and, for the aggregated volume, against the C restatement.  Shapes are
chosen around the kernel's own edges: heights below / at / just above one, two and three bands (32 rows by default, 16 in
the variant the last test runs in a subprocess), widths that are not
multiples of the 4-pixel chunks of the band pipeline, rows shorter than the prefetch distance, batches.
"""
import numpy as np
import pytest
import torch

import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair

pytestmark = pytest.mark.gpu


def _kw(D, mode, **over):
    kw = dict(minDisparity=0, numDisparities=D, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, preFilterCap=63,
              uniquenessRatio=15, speckleWindowSize=0, speckleRange=2, mode=mode)
    kw.update(over)
    return kw


CASES = []
for D in (128, 256):
    for h in (1, 2, 15, 16, 17, 31, 32, 33, 49, 63, 64, 65, 70, 97):
        for dw in (16, 17, 18, 19, 23, 37, 64, 131):
            if (h * 7 + dw) % 3 == 0 or h in (17, 33, 65):        # a third of the grid, every band edge (16- and 32-row bands)
                CASES.append((D, h, D + dw + 3, 1 if (h + dw) % 2 else 0))


@pytest.mark.parametrize("D,h,w,mode", CASES, ids=["D%d-h%d-w%d-m%d" % c for c in CASES])
def test_wave8_shapes_match_cv2(D, h, w, mode):
    import cv2
    kw = _kw(D, mode)
    l, r = synth_pair(w, h, D, seed=D + 3 * h + w)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    got = sm.StereoSGBM_create(**kw).compute(l, r)
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


@pytest.mark.parametrize("seed", range(24))
def test_wave8_random_parameters(seed):
    """P1 / P2 / uniqueness / disp12MaxDiff / blockSize / negative minDisparity sweeps inside the no-overflow domain."""
    import cv2
    rng = np.random.default_rng(4200 + seed)
    D = int(rng.choice([128, 128, 256]))
    bs = int(rng.choice([1, 3, 5]))
    cap = int(rng.choice([15, 31, 63]))
    cmax = bs * bs * (2 * (max(cap, 15) | 1) + 63)
    P2 = int(rng.integers(8, max(9, 8192 - cmax)))                 # 4 * (cmax + P2) <= 32768: the summed fast path
    P1 = int(rng.integers(1, P2))
    minD = int(rng.choice([0, 0, 5, -7, -40]))
    w = max(D + minD, 0) + int(rng.integers(20, 150)) + abs(min(minD, 0))
    h = int(rng.integers(1, 60))
    kw = _kw(D, int(rng.integers(0, 2)), blockSize=bs, preFilterCap=cap, P1=P1, P2=P2, minDisparity=minD,
             uniquenessRatio=int(rng.choice([0, 5, 15, 40, 99])), disp12MaxDiff=int(rng.choice([0, 1, 3, 1000])),
             speckleWindowSize=int(rng.choice([0, 50])))
    l, r = synth_pair(w, h, max(D + minD, 8), seed=500 + seed)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    m = sm.StereoSGBM_create(**kw)
    got = m.compute(l, r)
    assert not m.overflowed()
    assert np.array_equal(got, want), "%d pixels differ for %r %r" % ((got != want).sum(), kw, l.shape)


@pytest.mark.parametrize("D,h,w", [(128, 37, 190), (256, 21, 300)])
def test_wave8_volume_matches_oracle(D, h, w, port):
    """The aggregated S volume of the lean sweeps against the restatement (stage tap, every cell)."""
    kw = _kw(D, 1)
    l, r = synth_pair(w, h, D, seed=77)
    _, Cw, Sw, _ = port.sgbm_compute(l, r, want_volumes=True, **kw)
    m = sm.StereoSGBM_create(**kw)
    m.keep_volumes(True)
    got = m.compute(torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()).cpu().numpy()
    cost, S, g = m.debug_volumes(1, h, w)
    cost = cost[0, :, :, :D].cpu().numpy().astype(np.int32) & 0xffff
    S = S[0, :, :, :D].cpu().numpy()
    assert np.array_equal(cost + g["P2"], Cw.astype(np.int32)), "block cost differs"
    assert np.array_equal(S, Sw), "%d cells of S differ" % (S != Sw).sum()
    assert np.array_equal(got, port.sgbm_compute(l, r, **kw))


def test_wave8_batch_equals_singles_and_is_deterministic():
    """Many pairs in flight (bands of different pairs share SMs through the ticket order) == one pair at a time."""
    kw = _kw(128, 1, speckleWindowSize=100)
    pairs = [synth_pair(211, 53, 128, seed=s) for s in range(9)]
    L = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    R = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    m = sm.StereoSGBM_create(**kw)
    out = m.compute(L, R).cpu().numpy()
    again = m.compute(L, R).cpu().numpy()
    assert np.array_equal(out, again)
    single = sm.StereoSGBM_create(**kw)
    for i, (l, r) in enumerate(pairs):
        assert np.array_equal(out[i], single.compute(l, r)), "pair %d" % i


_ROWS16_WORKER = r"""
import sys
sys.path.insert(0, %r)
import numpy as np, cv2
import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair
ok = True
for (h, w, mode) in ((15, 160, 1), (16, 147, 1), (17, 190, 1), (33, 211, 0), (49, 150, 1), (70, 165, 1)):
    kw = dict(minDisparity=0, numDisparities=128, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, preFilterCap=63,
              uniquenessRatio=15, speckleWindowSize=0, speckleRange=2, mode=mode)
    l, r = synth_pair(w, h, 128, seed=h + w)
    ok = ok and np.array_equal(sm.StereoSGBM_create(**kw).compute(l, r), cv2.StereoSGBM_create(**kw).compute(l, r))
print("ROWS16_OK" if ok else "ROWS16_DIFFERS")
"""


def test_wave8_sixteen_row_variant(tmp_path):
    """B2S_WAVE8_ROWS=16 (128-thread CTAs, 16-row bands: the A/B variant) is read once per process: run it in its own."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "rows16.py"
    script.write_text(_ROWS16_WORKER % root)
    out = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=600,
                         env=dict(os.environ, B2S_WAVE8_ROWS="16"))
    assert "ROWS16_OK" in out.stdout, (out.stdout[-1000:], out.stderr[-2000:])


# ---- k_cost_fast2 (two columns per thread; D = 128, blockSize <= 5): strips of 64 - (blockSize - 1) output columns ----
COST2_CASES = [(w1, bs, minD, cn) for w1 in (3, 19, 55, 56, 59, 60, 61, 63, 64, 65, 119, 120, 121, 181) for bs, minD, cn in
               ((5, 0, 1), (3, 0, 1), (1, 0, 1))] + [(61, 5, -37, 1), (120, 3, 11, 1), (75, 5, 0, 3), (121, 3, -128, 3), (60, 1, 0, 3)]


COST2_CASES_256 = [(19, 5, 0, 1), (59, 5, 0, 1), (60, 3, 0, 1), (61, 5, 0, 1), (121, 1, 0, 1), (75, 5, -100, 1), (64, 3, 0, 3)]


@pytest.mark.parametrize("w1,bs,minD,cn", COST2_CASES_256, ids=["w%d-bs%d-m%d-c%d" % c for c in COST2_CASES_256])
def test_cost2_volume_matches_oracle_d256(w1, bs, minD, cn, port):
    """numDisparities = 256: the strip's disparities are split over two CTAs of the two-column kernel."""
    test_cost2_volume_matches_oracle(w1, bs, minD, cn, port, D=256)


@pytest.mark.parametrize("w1,bs,minD,cn", COST2_CASES, ids=["w%d-bs%d-m%d-c%d" % c for c in COST2_CASES])
def test_cost2_volume_matches_oracle(w1, bs, minD, cn, port, D=128):
    """Raw block cost C and aggregated S against the restatement around every strip edge of the two-column cost kernel
    (first / last strip clamp their window on the cropped axis), gray and colour, shifted and negative ranges."""
    from vectors import color_pair
    h = 13
    w = w1 + max(D + minD, 0) - min(minD, 0)
    assert w - max(D + minD, 0) + min(minD, 0) == w1
    kw = _kw(D, 1, blockSize=bs, minDisparity=minD, P1=8 * bs * bs, P2=32 * bs * bs)
    if w1 <= bs // 2:
        pytest.skip("cv2 rejects this width")
    l, r = (synth_pair(w, h, max(D + minD, 8), seed=w1 + bs) if cn == 1 else color_pair(w, h, max(D + minD, 8), w1 + bs))
    _, Cw, Sw, _ = port.sgbm_compute(l, r, want_volumes=True, **kw)
    m = sm.StereoSGBM_create(**kw)
    m.keep_volumes(True)
    got = m.compute(torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()).cpu().numpy()
    cost, S, g = m.debug_volumes(1, h, w)
    assert g["width1"] == w1
    cost = cost[0, :, :, :D].cpu().numpy().astype(np.int32) & 0xffff
    assert np.array_equal(cost + g["P2"], Cw.astype(np.int32)), "%d cost cells differ" % (cost + g["P2"] != Cw).sum()
    assert np.array_equal(S[0, :, :, :D].cpu().numpy(), Sw)
    assert np.array_equal(got, port.sgbm_compute(l, r, **kw))
