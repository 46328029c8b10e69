"""Device-resident throughput of the other BASELINE.json configurations (cfg1, cfg2, cfg4) next to cv2 on one
host core.  Informational: the judged line is bench.py (cfg3/cfg5).  usage: python tools/bench_configs.py [cfg1|cfg2|cfg4]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
import cv2
import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair
from vectors import CFG1, CFG2, CFG4


def timed(fn, iters):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def cpu_time(fn, reps=2):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e3


out = []
cv2.setNumThreads(1)
ONLY = sys.argv[1] if len(sys.argv) > 1 else ""
for name, cfg, B, D in (("cfg1_sgbm_640x480_D64", CFG1, 128, 64), ("cfg4_hh_2864x1988_D256", CFG4, 12, 256)):
    if ONLY and ONLY not in name:
        continue
    l, r = synth_pair(cfg["w"], cfg["h"], D, seed=cfg["seed"])
    L = torch.from_numpy(np.stack([l] * B)).cuda()
    R = torch.from_numpy(np.stack([r] * B)).cuda()
    m = sm.StereoSGBM_create(**cfg["kw"])
    ms = timed(lambda: m.compute(L, R), 3)
    ref = cv2.StereoSGBM_create(**cfg["kw"])
    want = ref.compute(l, r)
    ok = bool(np.array_equal(m.compute(L, R)[0].cpu().numpy(), want))
    cms = cpu_time(lambda: ref.compute(l, r), 1 if "cfg4" in name else 3)
    out.append({"config": name, "batch": B, "ms_per_batch": ms, "pairs_per_s": B / ms * 1e3,
                "gde_per_s": B / ms * 1e3 * cfg["w"] * cfg["h"] * D / 1e9, "cv2_1thread_ms_per_pair": cms, "parity_ok": ok})
    del m, L, R
    torch.cuda.empty_cache()
if ONLY and ONLY != "cfg2":
    for o in out:
        print(json.dumps(o))
    sys.exit(0)
l, r = synth_pair(CFG2["w"], CFG2["h"], CFG2["D"], seed=CFG2["seed"])
B = 64
L = torch.from_numpy(np.stack([l] * B)).cuda()
R = torch.from_numpy(np.stack([r] * B)).cuda()
m = sm.StereoBM_create(numDisparities=CFG2["D"], blockSize=CFG2["bs"])
ms = timed(lambda: m.compute(L, R), 3)
ref = cv2.StereoBM_create(numDisparities=CFG2["D"], blockSize=CFG2["bs"])
ok = bool(np.array_equal(m.compute(L, R)[0].cpu().numpy(), ref.compute(l, r)))
cms = cpu_time(lambda: ref.compute(l, r), 3)
out.append({"config": "cfg2_bm_1280x720_D128_bs15", "batch": B, "ms_per_batch": ms, "pairs_per_s": B / ms * 1e3,
            "gde_per_s": B / ms * 1e3 * 1280 * 720 * 128 / 1e9, "cv2_1thread_ms_per_pair": cms, "parity_ok": ok})
for o in out:
    print(json.dumps(o))
