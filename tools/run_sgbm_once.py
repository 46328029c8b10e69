"""Runs the cfg3 matcher a few times on a batch (for ncu captures).  usage: run_sgbm_once.py [batch] [iters] [mode]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
it = int(sys.argv[2]) if len(sys.argv) > 2 else 2
mode = int(sys.argv[3]) if len(sys.argv) > 3 else 1
ps = [synth_pair(1242, 375, 128, seed=s) for s in range(4)]
L = torch.from_numpy(np.stack([ps[i % 4][0] for i in range(B)])).cuda()
R = torch.from_numpy(np.stack([ps[i % 4][1] for i in range(B)])).cuda()
m = sm.StereoSGBM_create(minDisparity=0, numDisparities=128, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1,
                         uniquenessRatio=15, speckleWindowSize=100, speckleRange=2, preFilterCap=63, mode=mode)
m.enable_timing(True)
for _ in range(it):
    d = m.compute(L, R)
    torch.cuda.synchronize()
    print(m.stage_ms())
