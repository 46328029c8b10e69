import sys; sys.path.insert(0,'/root/repo')
import numpy as np, torch, time
import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair
ps=[synth_pair(1242,375,128,seed=s) for s in range(2)]
for B in (1,8,32):
    L=torch.from_numpy(np.stack([ps[i%2][0] for i in range(B)])).cuda(); R=torch.from_numpy(np.stack([ps[i%2][1] for i in range(B)])).cuda()
    m=sm.StereoSGBM_create(minDisparity=0,numDisparities=128,blockSize=5,P1=600,P2=2400,disp12MaxDiff=1,uniquenessRatio=15,speckleWindowSize=100,speckleRange=2,preFilterCap=63,mode=1)
    m.enable_timing(True)
    for _ in range(3): m.compute(L,R)
    torch.cuda.synchronize()
    st=m.stage_ms(); sw=m.sweep_ms()
    t0=time.perf_counter()
    for _ in range(20): m.compute(L,R)
    torch.cuda.synchronize()
    wall=(time.perf_counter()-t0)/20*1e3
    print(B, 'wall %.2f ms'%wall, {k:round(v,3) for k,v in st.items()}, 'sweeps', [round(x,3) for x in sw])
