#!/usr/bin/env python
"""bench.py -- throughput of the stereo hot path on B200 (and of the reference's CPU path).

Workload (BASELINE.json metric): synthetic rectifiable KITTI-shape
pairs, 1242x375, per pair
    remap(left), remap(right)  ->  StereoSGBM MODE_HH D=128 bs=5 P1=600 P2=2400 LR=1 uniq=15
    speckle(100,2)  ->  reprojectImageTo3D(disp/16, Q)
A "step" pushes one batch of B pairs through that chain.  One process per GPU, pairs sharded
per rank, no collective on the data path (weak scaling: B pairs per GPU per step).

  value : pairs/s, whole job, inputs already resident in HBM, CUDA-event timed, max over ranks
  e2e   : same chain through the public Python API with HOST (pinned) buffers: H2D of the raw
          pair, D2H of the int16 disparity and the float32 XYZ inside the timed region
  roofline     : the path-aggregation stage (the two k_wave sweeps per step) against measured HBM peak
  cpu_baseline : cv2 4.13.0 (the reference's own implementation) on this box's host cores

`--impl reference` times cv2 itself (all host cores, pair-parallel worker processes).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, D = 1242, 375, 128
SGBM_KW = dict(minDisparity=0, numDisparities=D, blockSize=5, P1=600, P2=2400, disp12MaxDiff=1,
               uniquenessRatio=15, speckleWindowSize=100, speckleRange=2, preFilterCap=63, mode=1)
GDE_PER_PAIR = W * H * D / 1e9
WORKLOAD = "kitti_1242x375_D128_MODE_HH_remap+sgbm+speckle+reproject (BASELINE configs[2]/[4])"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_pairs(n, seed0=0):
    """n synthetic pairs (uint8 [n,H,W] x2); a few distinct pairs tiled to n to bound setup time."""
    from stereo_match_b200.synth import synth_pair
    distinct = min(n, 8)
    ps = [synth_pair(W, H, D, seed=seed0 + s) for s in range(distinct)]
    L = np.stack([ps[i % distinct][0] for i in range(n)])
    R = np.stack([ps[i % distinct][1] for i in range(n)])
    return L, R


# --------------------------------------------------------------------------------------
# reference arm / cpu_baseline: cv2 on the host cores, pair-parallel worker processes
# --------------------------------------------------------------------------------------
def _cv2_worker(args):
    npairs, seed = args
    import cv2
    from stereo_match_b200.synth import synth_pair, KITTI_Q, kitti_rect_maps
    cv2.setNumThreads(1)
    l, r = synth_pair(W, H, D, seed=seed)
    mx, my = kitti_rect_maps(W, H)
    m = cv2.StereoSGBM_create(**SGBM_KW)
    t0 = time.perf_counter()
    for _ in range(npairs):
        lr = cv2.remap(l, mx, my, cv2.INTER_LINEAR)
        rr = cv2.remap(r, mx, my, cv2.INTER_LINEAR)
        d = m.compute(lr, rr)
        cv2.reprojectImageTo3D(d.astype(np.float32) / 16.0, KITTI_Q.astype(np.float32))
    return time.perf_counter() - t0


def cv2_pairs_per_s(pool, cores, pairs_per_worker):
    t0 = time.perf_counter()
    pool.map(_cv2_worker, [(pairs_per_worker, 1000 + i) for i in range(cores)])
    dt = time.perf_counter() - t0
    return cores * pairs_per_worker / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    import cv2
    cores = os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        for _ in range(args.warmup):
            cv2_pairs_per_s(pool, cores, 1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            cv2_pairs_per_s(pool, cores, 1)
        dt = time.perf_counter() - t0
    pairs = cores * args.steps
    v = pairs / dt
    line = {
        "impl": "reference", "metric": "pairs_per_s", "value": v, "unit": "pairs/s", "gde_per_s": v * GDE_PER_PAIR,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_per_step": cores, "note": "cv2 %s, %d worker processes x cv2.setNumThreads(1)" % (cv2.__version__, cores)},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "reference",
                         "sample": "%d steps x %d pairs (one per worker process), remap x2 + StereoSGBM MODE_HH + reprojectImageTo3D" % (args.steps, cores)},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# --------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# --------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import stereo_match_b200 as sm
    from stereo_match_b200 import _lib
    from stereo_match_b200.synth import KITTI_Q, kitti_rect_maps

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("WORLD_SIZE=%d but --gpus %d" % (world, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch

    # ---- inputs: rank r owns pairs [r*B, (r+1)*B) of the job; resident in HBM --------------
    Lh, Rh = make_pairs(B, seed0=rank * 8)
    mx, my = kitti_rect_maps(W, H)
    Lh, Rh = torch.from_numpy(Lh).pin_memory(), torch.from_numpy(Rh).pin_memory()
    Ld, Rd = Lh.to(dev), Rh.to(dev)
    mxd, myd = torch.from_numpy(mx).to(dev), torch.from_numpy(my).to(dev)
    Q = KITTI_Q.astype(np.float32)
    matcher = sm.StereoSGBM_create(**SGBM_KW)
    matcher.enable_timing(True)

    def chain(Lraw, Rraw, m=None):
        Lr = sm.remap(Lraw.unsqueeze(-1), mxd, myd, sm.INTER_LINEAR).squeeze(-1)
        Rr = sm.remap(Rraw.unsqueeze(-1), mxd, myd, sm.INTER_LINEAR).squeeze(-1)
        disp = (m or matcher).compute(Lr, Rr)
        xyz = sm.reprojectImageTo3D(disp, Q, disp_scale=1.0 / 16.0)
        return disp, xyz

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    for _ in range(args.warmup):
        chain(Ld, Rd)
    barrier()

    # ---- timed region 1: resident inputs ------------------------------------------------
    lib = _lib.load()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    stage_sum = {}
    launches0 = lib.b2s_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        chain(Ld, Rd)
        for k, v in matcher.stage_ms().items():       # waits for this step's last stage event
            stage_sum[k] = stage_sum.get(k, 0.0) + v
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = int(lib.b2s_launch_count() - launches0)
    clocks = sampler.stop() if rank == 0 else None
    value = world * B * args.steps / (ms_total / 1e3)

    # ---- timed region 2: end to end with host buffers -------------------------------------
    # The call a user makes (remap -> compute -> reprojectImageTo3D) on HOST data: every step copies its
    # raw pair from pinned memory and reads disparity + XYZ back.  Two CUDA streams (one matcher handle
    # each, handles own their workspace) alternate steps so the copies of one step overlap the kernels
    # of the other; nothing is skipped or cached.
    NS = args.e2e_streams
    streams = [torch.cuda.Stream(device=dev) for _ in range(NS)]
    matchers = [matcher] + [sm.StereoSGBM_create(**SGBM_KW) for _ in range(NS - 1)]
    disp_h = [torch.empty((B, H, W), dtype=torch.int16).pin_memory() for _ in range(NS)]
    xyz_h = [torch.empty((B, H, W, 3), dtype=torch.float32).pin_memory() for _ in range(NS)]
    Ls = [torch.empty_like(Ld) for _ in range(NS)]
    Rs = [torch.empty_like(Rd) for _ in range(NS)]
    keep = [None] * NS

    def e2e_step(i):
        k = i % NS
        with torch.cuda.stream(streams[k]):
            Ls[k].copy_(Lh, non_blocking=True)
            Rs[k].copy_(Rh, non_blocking=True)
            disp, xyz = chain(Ls[k], Rs[k], matchers[k])
            disp_h[k].copy_(disp, non_blocking=True)
            xyz_h[k].copy_(xyz, non_blocking=True)
            keep[k] = (disp, xyz)               # device buffers stay alive until the stream has drained

    def e2e_sync():
        for st in streams:
            st.synchronize()

    matcher.enable_timing(False)
    for i in range(max(NS, args.warmup)):
        e2e_step(i)
    e2e_sync()
    barrier()
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        e2e_step(i)
    for st in streams:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    wall_e2e = (time.perf_counter() - t0) * 1e3
    e2e_value = world * B * args.steps / (ms_e2e / 1e3)
    h2d = int(Lh.numel() + Rh.numel())
    d2h = int(disp_h[0].numel() * 2 + xyz_h[0].numel() * 4)
    disp_h, xyz_h = disp_h[(args.steps - 1) % NS], xyz_h[(args.steps - 1) % NS]

    # ---- parity gate of the run: the result just produced against cv2 on pair 0 ------------
    parity = None
    if rank == 0 and not args.no_parity:
        import cv2
        lr = cv2.remap(Lh[0].numpy(), mx, my, cv2.INTER_LINEAR)
        rr = cv2.remap(Rh[0].numpy(), mx, my, cv2.INTER_LINEAR)
        want = cv2.StereoSGBM_create(**SGBM_KW).compute(lr, rr)
        wxyz = cv2.reprojectImageTo3D(want.astype(np.float32) / 16.0, Q)
        parity = bool(np.array_equal(disp_h[0].numpy(), want) and
                      np.array_equal(xyz_h[0].numpy().view(np.uint32), wxyz.view(np.uint32)))

    # ---- roofline of the dominant stage -----------------------------------------------------
    g = matcher.geometry(W)
    agg_ms = stage_sum["aggregate"] / args.steps
    alg_bytes_step = 8.0 * g["width1"] * H * D * B          # C w+r, S w+r, int16
    peak, peak_src = measured_peaks()
    achieved = alg_bytes_step / (agg_ms / 1e3) / 1e9
    nl = 2                                                   # k_wave launches per step: forward + mirrored sweep
    traffic, traffic_src = args.traffic, "command line" if args.traffic else None
    if traffic is None:
        # dram bytes per k_wave launch from the committed full ncu capture (tools/make_profiles.py);
        # only valid for the batch it was captured at
        import glob
        for fn in sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic_r*.json")))[::-1]:
            t = json.load(open(fn))
            if t.get("batch") == B:
                traffic, traffic_src = t["avg_dram_bytes_per_launch"], "profiles/" + os.path.basename(fn)
                break
    roofline = {"bound": "hbm", "kernel": "k_wave (2 launches per step: forward sweep of 4 path directions, mirrored sweep of the "
                                          "other 4 with the winner-take-all fused)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes_step / nl,
                "avg_launch_ms": agg_ms / nl, "traffic": traffic, "traffic_source": traffic_src,
                "stage_ms_per_step": {k: v / args.steps for k, v in stage_sum.items()},
                "stage_share_of_step": {k: v / args.steps / (ms_total / args.steps) for k, v in stage_sum.items()}}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import multiprocessing as mp
        cores = os.cpu_count() or 1
        with mp.get_context("spawn").Pool(cores) as pool:
            cv2_pairs_per_s(pool, cores, 1)                      # warm-up (imports, page-in)
            v, dt = cv2_pairs_per_s(pool, cores, args.cpu_pairs)
        cpu = {"value": v, "unit": "pairs/s", "cores": cores, "kind": "reference",
               "sample": "%d pairs (%d per worker process, cv2.setNumThreads(1) each), same chain, %.1f s" % (cores * args.cpu_pairs, args.cpu_pairs, dt)}

    if rank == 0:
        line = {
            "metric": "pairs_per_s", "value": value, "unit": "pairs/s", "gde_per_s": value * GDE_PER_PAIR,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_step_per_gpu": B, "parallelism": "pair-parallel x%d, no collective" % world,
                       "l2": "per-step working set %.1f GB >> 126 MB L2 (no flush needed)" % (matcher.workspace_bytes(H, W, B) / 1e9)},
            "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "wall_ms_per_step": wall_e2e / args.steps,
                    "how": "%d streams x 1 matcher handle each" % NS + ", pinned host buffers, H2D + kernels + D2H per step"},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity_ok": parity,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the real stdout; everything else printed while the run was going on
    (NCCL's version banner, library chatter) was diverted to stderr by quiet_stdout()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def quiet_stdout():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=128, help="pairs per step per GPU")
    ap.add_argument("--cpu-pairs", type=int, default=40, help="cpu_baseline: pairs per worker process")
    ap.add_argument("--e2e-streams", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--traffic", type=float, default=None,
                    help="dram bytes per k_aggregate launch from the committed ncu capture (profiles/)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus > 1 and "RANK" not in os.environ:
        raise SystemExit("launch with: python -m torch.distributed.run --nnodes=1 --nproc-per-node %d bench.py --gpus %d ..." % (args.gpus, args.gpus))
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
