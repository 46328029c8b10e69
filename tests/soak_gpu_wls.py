"""Randomised soak of the CUDA disparity WLS filter (stereo_match_b200/csrc/wls.cu) against its restatement
oracle/wls_oracle.c: bit-exact int16 output and bitwise confidence maps over random lambda / sigma / radius / LRC
threshold / ROI / sizes down to one row or one column / arbitrary int16 disparity maps / gray and colour guides.
(The restatement itself is PARITY UNPINNED against cv2.ximgproc, which exists on neither box.)  A script, not collected
by pytest:      python tests/soak_gpu_wls.py [n_cases] [seed0]      (GPU box)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import stereo_match_b200 as sm      # noqa: E402
from oracle import port             # noqa: E402

n, seed0 = (int(sys.argv[1]) if len(sys.argv) > 1 else 300), (int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for k in range(seed0, seed0 + n):
    rng = np.random.default_rng(272700 + k)
    H = int(rng.choice([1, 2, 3, rng.integers(4, 40), rng.integers(40, 160)]))
    W = int(rng.choice([1, 2, 3, rng.integers(4, 60), rng.integers(60, 260)]))
    cn = int(rng.choice([1, 3]))
    guide = rng.integers(0, 256, (H, W) if cn == 1 else (H, W, 3), dtype=np.uint8)
    if rng.random() < 0.5:                              # smooth guide with edges instead of noise
        yy, xx = np.mgrid[0:H, 0:W]
        g = ((xx // max(W // 5, 1) * 37 + yy // max(H // 4, 1) * 53) % 256).astype(np.uint8)
        guide = g if cn == 1 else np.stack([g, 255 - g, g // 2], -1)
    kind = int(rng.integers(0, 3))
    if kind == 0:                                       # disparity-like maps with holes
        dl = (rng.integers(0, 64, (H, W)) * 16 + rng.integers(0, 16, (H, W))).astype(np.int16)
        dr = (-(rng.integers(0, 64, (H, W)) * 16 + rng.integers(0, 16, (H, W)))).astype(np.int16)
        dl[rng.random((H, W)) < 0.3] = -16
        dr[rng.random((H, W)) < 0.3] = 16 * 64
    elif kind == 1:                                     # consistent pair: dr(x - dl/16) = -dl
        d = int(rng.integers(1, 40))
        dl = np.full((H, W), 16 * d, np.int16)
        dr = np.full((H, W), -16 * d, np.int16)
        dl[:, :min(d, W)] = -16
    else:                                               # anything an int16 can hold
        dl = rng.integers(-32768, 32768, (H, W)).astype(np.int16)
        dr = rng.integers(-32768, 32768, (H, W)).astype(np.int16)
    conf = bool(rng.random() < 0.7)
    x0 = int(rng.integers(0, W)); y0 = int(rng.integers(0, H))
    roi = (x0, y0, int(rng.integers(1, W - x0 + 1)), int(rng.integers(1, H - y0 + 1)))
    lam = float(rng.choice([0.0, 1.0, 80.0, 8000.0, 80000.0, 1e7]) * rng.uniform(0.5, 2.0))
    sigma = float(rng.choice([0.05, 0.5, 1.2, 3.0, 20.0]))
    radius = int(rng.choice([0, 1, 2, 5, 9, 30]))
    lrc = int(rng.choice([0, 1, 24, 100, 5000]))
    f = sm.createDisparityWLSFilterGeneric(conf)
    f.setLambda(lam); f.setSigmaColor(sigma); f.setDepthDiscontinuityRadius(radius); f.setLRCthresh(lrc)
    try:
        got = f.filter(dl, guide, None, dr if conf else None, roi)
    except sm.error as ex:
        try:
            port.wls_filter(dl, dr if conf else None, guide, roi, use_confidence=conf, lambda_=lam, sigma=sigma, radius=radius, lrc_thresh=lrc)
        except ValueError:
            continue
        bad += 1
        print("FAIL wls case %d: the engine raises (%s), the restatement does not" % (k, str(ex)[:90]))
        continue
    want, wconf = port.wls_filter(dl, dr if conf else None, guide, roi, use_confidence=conf, lambda_=lam, sigma=sigma,
                                  radius=radius, lrc_thresh=lrc)
    ok = np.array_equal(got, want)
    if ok and conf:
        ok = np.array_equal(f.getConfidenceMap().view(np.uint32), wconf.view(np.uint32))
    if not ok:
        bad += 1
        print("FAIL wls case %d: %d pixels differ (max %d), H %d W %d cn %d kind %d conf %d roi %r lambda %g sigma %g radius %d lrc %d"
              % (k, int((got != want).sum()), int(np.abs(got.astype(int) - want).max()), H, W, cn, kind, conf, roi, lam, sigma, radius, lrc))
print("soak wls: %d cases, %d failures" % (n, bad))
sys.exit(1 if bad else 0)
