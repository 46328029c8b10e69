"""Multi-GPU tests (skipped on a one-GPU box; run with `gpurun --gpus 2`): the optional NCCL gather of per-rank
results (north_star: "NCCL is used only to gather point clouds if requested"; stereo_match_b200/batch.py) on real
CUDA tensors, and one process driving two devices (the dynamic-shared-memory opt-in is per device)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_NCCL_WORKER = r"""
import os, sys, json
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
import stereo_match_b200 as sm
from stereo_match_b200.batch import shard_range, gather_host_results
from stereo_match_b200.synth import synth_pair, KITTI_Q
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
n = 5
lo, hi = shard_range(n, rank, world)
kw = dict(minDisparity=0, numDisparities=32, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, uniquenessRatio=15, preFilterCap=63, mode=1)
pairs = [synth_pair(150, 40, 32, seed=s) for s in range(n)]
L = torch.from_numpy(np.stack([pairs[i][0] for i in range(lo, hi)])).cuda()
R = torch.from_numpy(np.stack([pairs[i][1] for i in range(lo, hi)])).cuda()
m = sm.StereoSGBM_create(**kw)
disp = m.compute(L, R)
Q = KITTI_Q.astype(np.float32)
xyz = sm.reprojectImageTo3D(disp, Q, disp_scale=1.0 / 16.0)
alld = gather_host_results(disp, n, dst=0)            # CUDA tensors: ncclAllGather
allx = gather_host_results(xyz, n, dst=0)
if rank == 0:
    import cv2
    ok = alld.is_cuda and tuple(alld.shape) == (n, 40, 150) and tuple(allx.shape) == (n, 40, 150, 3)
    for i in range(n):
        want = cv2.StereoSGBM_create(**kw).compute(*pairs[i])
        ok = ok and np.array_equal(alld[i].cpu().numpy(), want)
        ok = ok and np.array_equal(allx[i].cpu().numpy().view(np.uint32),
                                   cv2.reprojectImageTo3D(want.astype(np.float32) / 16.0, Q).view(np.uint32))
    print(json.dumps({"ok": bool(ok)}))
dist.destroy_process_group()
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_nccl_gather_of_disparity_and_clouds(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_NCCL_WORKER % ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29617", str(script)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert lines and json.loads(lines[-1])["ok"], (out.stdout[-2000:], out.stderr[-2000:])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_two_devices():
    """Kernels that opt in to > 48 KB dynamic shared memory must work on every device of the process."""
    import cv2
    import stereo_match_b200 as sm
    from stereo_match_b200.synth import synth_pair
    kw = dict(minDisparity=0, numDisparities=128, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1, uniquenessRatio=15,
              preFilterCap=63, mode=1)
    l, r = synth_pair(260, 40, 128, seed=2)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    wbm = cv2.StereoBM_create(64, 21).compute(l, r)
    for d in (0, 1, 0):
        with torch.cuda.device(d):
            L, R = torch.from_numpy(l).cuda(d), torch.from_numpy(r).cuda(d)
            got = sm.StereoSGBM_create(**kw).compute(L, R)
            assert got.device.index == d and np.array_equal(got.cpu().numpy(), want)
            assert np.array_equal(sm.StereoBM_create(64, 21).compute(L, R).cpu().numpy(), wbm)
