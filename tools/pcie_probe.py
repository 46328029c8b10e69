"""Host-link ceiling for the e2e numbers of bench.py: pinned-memory D2H, H2D and both at once, on 1 GPU or on N GPUs
CONCURRENTLY (one process per GPU, all copying between the same two barriers), for a few transfer sizes.

    python tools/pcie_probe.py                                        # this GPU alone
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py

Rank 0 prints one JSON line: per size and direction the per-GPU rate (min over ranks) and the aggregate over all GPUs.
bench.py's e2e moves 119 MB up and 835 MB down per step and GPU; `value_e2e <= aggregate_d2h / 6.52 MB per pair`.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def reduce(v, op):
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=op)
    return float(t.item())


out = {"n_gpus": world, "sizes": {}}
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
for mb in (64, 256, 835):
    n = mb * 1000 * 1000
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    d2 = torch.empty(n, dtype=torch.uint8, device=dev)
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
    res = {}
    for name in ("d2h", "h2d", "both"):
        reps = 6

        def go():
            if name in ("d2h", "both"):
                with torch.cuda.stream(s_dn):
                    h.copy_(d, non_blocking=True)
            if name in ("h2d", "both"):
                with torch.cuda.stream(s_up):
                    d2.copy_(h2, non_blocking=True)

        go()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            go()
        torch.cuda.current_stream().wait_stream(s_dn)
        torch.cuda.current_stream().wait_stream(s_up)
        e1.record()
        barrier()
        ms = reduce(e0.elapsed_time(e1), dist.ReduceOp.MAX if world > 1 else None)
        per_dir = reps * n / ms / 1e6                       # GB/s per GPU and direction (slowest rank)
        res[name] = {"GBps_per_gpu_per_direction": per_dir, "aggregate_GBps_per_direction": per_dir * world}
    out["sizes"]["%d_MB" % mb] = res
    del d, d2, h, h2
if rank == 0:
    sys.stdout.write(json.dumps(out) + "\n")
if world > 1:
    dist.destroy_process_group()
