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
WORKLOAD = ("kitti_1242x375_D128_MODE_HH_remap+sgbm+speckle+reproject (BASELINE configs[2]/[4]; the left-view chain only: "
            "the right-view matcher and the WLS filter of compute_disparity are not part of the BASELINE metric)")


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
# reference arm / cpu_baseline: cv2 on the host cores, pair-parallel PERSISTENT worker processes
# --------------------------------------------------------------------------------------
def _cv2_worker(idx, k, nwarm, nsteps, barrier, out):
    """One worker = one host core.  Everything that is not the per-pair chain -- synthetic input, rectification maps,
    matcher object -- is built ONCE, before the timed region; then `nwarm` untimed and `nsteps` timed steps of k pairs."""
    import cv2
    from stereo_match_b200.synth import synth_pair, KITTI_Q, kitti_rect_maps
    cv2.setNumThreads(1)
    l, r = synth_pair(W, H, D, seed=1000 + idx)
    mx, my = kitti_rect_maps(W, H)
    m = cv2.StereoSGBM_create(**SGBM_KW)
    Q = KITTI_Q.astype(np.float32)

    def one_pair():
        lr = cv2.remap(l, mx, my, cv2.INTER_LINEAR)
        rr = cv2.remap(r, mx, my, cv2.INTER_LINEAR)
        d = m.compute(lr, rr)
        return cv2.reprojectImageTo3D(d.astype(np.float32) / 16.0, Q)

    for _ in range(nwarm * k):
        one_pair()
    barrier.wait()                       # start of the timed region (the parent takes the time at the same barrier)
    t0 = time.perf_counter()
    for _ in range(nsteps * k):
        one_pair()
    out[idx] = time.perf_counter() - t0
    barrier.wait()                       # end of the timed region


def cv2_pairs_per_s(cores, k, nwarm, nsteps):
    """pairs/s of cv2 over `cores` worker processes x cv2.setNumThreads(1) (StereoSGBM MODE_HH is serial inside OpenCV,
    SURVEY.md section 6): cores * nsteps * k pairs between two barriers."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    barrier = ctx.Barrier(cores + 1)
    out = ctx.Array("d", cores)
    procs = [ctx.Process(target=_cv2_worker, args=(i, k, nwarm, nsteps, barrier, out), daemon=True) for i in range(cores)]
    for p in procs:
        p.start()
    barrier.wait()
    t0 = time.perf_counter()
    barrier.wait()
    dt = time.perf_counter() - t0
    for p in procs:
        p.join()
    return cores * nsteps * k / dt, dt, max(out[:])


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import cv2
    cores = os.cpu_count() or 1
    k = args.ref_pairs
    v, dt, slowest = cv2_pairs_per_s(cores, k, max(args.warmup, 1), args.steps)      # W untimed steps, then K timed ones
    line = {
        "impl": "reference", "metric": "pairs_per_s", "value": v, "unit": "pairs/s", "gde_per_s": v * GDE_PER_PAIR,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_per_step": cores * k,
                   "note": "cv2 %s, %d persistent worker processes x cv2.setNumThreads(1); inputs, maps and matcher built "
                           "once per worker before the timed region" % (cv2.__version__, cores)},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "reference",
                         "sample": "%d steps x %d pairs (%d per worker process), remap x2 + StereoSGBM MODE_HH + speckle + "
                                   "reprojectImageTo3D, %.1f s wall (slowest worker %.1f s)" % (args.steps, cores * k, k, dt, slowest)},
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
# cfg5 job, other configurations, latency
# --------------------------------------------------------------------------------------
def _synth_job_pair(seed):
    from stereo_match_b200.synth import synth_pair
    return synth_pair(W, H, D, seed=seed)


def run_job(args, sm, torch, dist, dev, rank, world, chain, matcher, barrier, max_over_ranks, mx, my, Q):
    """BASELINE configs[4]: a batch of `--job` synthetic KITTI-shape pairs (seeds 0 .. N-1, all distinct),
    remap + SGBM + reprojection, sharded pair-parallel: rank r owns the contiguous block shard_range(N, r, world),
    streams it from pinned host memory in batches of B, reads disparity and XYZ back, and checksums the XYZ bits of
    every pair on the device.  One pair per rank (a different one on every rank) is checked against cv2."""
    import multiprocessing as mp
    from stereo_match_b200.batch import shard_range
    B = args.batch
    lo, hi = shard_range(args.job, rank, world)
    n = hi - lo
    with mp.get_context("spawn").Pool(max(1, (os.cpu_count() or 1) // world)) as pool:
        pairs = pool.map(_synth_job_pair, range(lo, hi), chunksize=8)
    Lall = torch.from_numpy(np.stack([p[0] for p in pairs])).pin_memory()
    Rall = torch.from_numpy(np.stack([p[1] for p in pairs])).pin_memory()
    del pairs
    NS = 2
    streams = [torch.cuda.Stream(device=dev) for _ in range(NS)]
    ms = [matcher, sm.StereoSGBM_create(**SGBM_KW)]
    Ls = [torch.empty((B, H, W), dtype=torch.uint8, device=dev) for _ in range(NS)]
    Rs = [torch.empty((B, H, W), dtype=torch.uint8, device=dev) for _ in range(NS)]
    disp_h = [torch.empty((B, H, W), dtype=torch.int16).pin_memory() for _ in range(NS)]
    xyz_h = [torch.empty((B, H, W, 3), dtype=torch.float32).pin_memory() for _ in range(NS)]
    sums = torch.zeros(n, dtype=torch.int64, device=dev)
    pick = lo + (7919 * (rank + 1)) % n                       # the pair this rank checks against cv2
    keep_pick = {}
    keep = [None] * NS

    def run(measure):
        for bi, b0 in enumerate(range(0, n, B)):
            k, nb = bi % NS, min(B, n - b0)
            with torch.cuda.stream(streams[k]):
                Ls[k][:nb].copy_(Lall[b0:b0 + nb], non_blocking=True)
                Rs[k][:nb].copy_(Rall[b0:b0 + nb], non_blocking=True)
                disp, xyz = chain(Ls[k][:nb], Rs[k][:nb], ms[k])
                disp_h[k][:nb].copy_(disp, non_blocking=True)
                xyz_h[k][:nb].copy_(xyz, non_blocking=True)
                sums[b0:b0 + nb] = xyz.view(torch.int32).to(torch.int64).sum(dim=(1, 2, 3))
                if measure and b0 <= pick - lo < b0 + nb:
                    keep_pick["d"] = disp[pick - lo - b0].clone()
                    keep_pick["x"] = xyz[pick - lo - b0].clone()
                keep[k] = (disp, xyz)
        for st in streams:
            st.synchronize()

    run(False)                                                # warm-up pass (workspaces, page-in)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run(True)
    e1.record()
    barrier()
    ms_job = max_over_ranks(e0.elapsed_time(e1))
    import cv2
    cv2.setNumThreads(max(1, (os.cpu_count() or 1) // world))
    lr = cv2.remap(Lall[pick - lo].numpy(), mx, my, cv2.INTER_LINEAR)
    rr = cv2.remap(Rall[pick - lo].numpy(), mx, my, cv2.INTER_LINEAR)
    want = cv2.StereoSGBM_create(**SGBM_KW).compute(lr, rr)
    wxyz = cv2.reprojectImageTo3D(want.astype(np.float32) / 16.0, Q)
    ok = bool(np.array_equal(keep_pick["d"].cpu().numpy(), want) and
              np.array_equal(keep_pick["x"].cpu().numpy().view(np.uint32), wxyz.view(np.uint32)))
    ok = ok and int(sums[pick - lo].item()) == int(wxyz.view(np.int32).astype(np.int64).sum())
    info = torch.tensor([1 if ok else 0, pick, int(sums.sum().item() & 0x7fffffffffffffff)], dtype=torch.int64, device=dev)
    if world > 1:
        allinfo = [torch.zeros_like(info) for _ in range(world)]
        dist.all_gather(allinfo, info)
    else:
        allinfo = [info]
    return {"pairs": args.job, "distinct": True, "seeds": "0..%d" % (args.job - 1), "batch": B, "shard": "contiguous (batch.shard_range)",
            "seconds": ms_job / 1e3, "pairs_per_s": args.job / (ms_job / 1e3),
            "h2d_bytes": int(args.job) * 2 * H * W, "d2h_bytes": int(args.job) * H * W * 14,
            "parity_ok_per_rank": [bool(int(t[0].item())) for t in allinfo], "pair_checked_per_rank": [int(t[1].item()) for t in allinfo],
            "xyz_checksum_per_rank": [int(t[2].item()) for t in allinfo]}


def run_extras(sm, torch, chain, Ld, Rd, Lh, mx, my, Q, peak):
    """cfg1 / cfg2 / cfg4 of BASELINE.json (device-resident, parity against cv2 on one pair, dominant-stage roofline)
    and the cfg3 chain at batch 1 and 8 (a video frame at a time)."""
    import cv2
    from stereo_match_b200.synth import synth_pair
    cv2.setNumThreads(os.cpu_count() or 1)

    def timed(fn, iters):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters

    other = {}
    cfgs = (("cfg1_sgbm_640x480_D64_bs5", 640, 480, 64, 1, 128, dict(SGBM_KW, numDisparities=64, mode=0, speckleWindowSize=0)),
            ("cfg4_hh_2864x1988_D256", 2864, 1988, 256, 4, 8, dict(SGBM_KW, numDisparities=256)))
    for name, w, h, d, seed, nb, kw in cfgs:
        l, r = synth_pair(w, h, d, seed=seed)
        L = torch.from_numpy(np.stack([l] * nb)).cuda()
        R = torch.from_numpy(np.stack([r] * nb)).cuda()
        L[-1], R[-1] = torch.from_numpy(l[::-1].copy()).cuda(), torch.from_numpy(r[::-1].copy()).cuda()   # not all pairs alike
        m = sm.StereoSGBM_create(**kw)
        m.enable_timing(True)
        ms = timed(lambda: m.compute(L, R), 3)
        st = m.stage_ms()
        out = m.compute(L, R)
        ok = bool(np.array_equal(out[0].cpu().numpy(), cv2.StereoSGBM_create(**kw).compute(l, r)) and
                  np.array_equal(out[-1].cpu().numpy(), cv2.StereoSGBM_create(**kw).compute(l[::-1].copy(), r[::-1].copy())))
        g = m.geometry(w)
        ndir = g["ndirs"]
        # algorithmic bytes of the aggregation stage: MODE_HH C r+r, S w+r; MODE_SGBM has no S1 round trip
        alg = (8.0 if ndir == 8 else 4.0) * g["width1"] * h * d * nb
        ach = alg / (st["aggregate"] / 1e3) / 1e9
        other[name] = {"batch": nb, "ms_per_batch": ms, "pairs_per_s": nb / ms * 1e3, "gde_per_s": nb / ms * 1e3 * w * h * d / 1e9,
                       "parity_ok": ok, "stage_ms": st,
                       "roofline": {"kernel": "aggregation sweeps", "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                                    "frac": ach / peak, "algorithmic_bytes": alg}}
        del m, L, R, out
        torch.cuda.empty_cache()
    l, r = synth_pair(1280, 720, 128, seed=2)
    nb = 64
    L = torch.from_numpy(np.stack([l] * nb)).cuda()
    R = torch.from_numpy(np.stack([r] * nb)).cuda()
    bm = sm.StereoBM_create(numDisparities=128, blockSize=15)
    ms = timed(lambda: bm.compute(L, R), 3)
    ok = bool(np.array_equal(bm.compute(L, R)[nb - 1].cpu().numpy(), cv2.StereoBM_create(numDisparities=128, blockSize=15).compute(l, r)))
    other["cfg2_bm_1280x720_D128_bs15"] = {"batch": nb, "ms_per_batch": ms, "pairs_per_s": nb / ms * 1e3,
                                           "gde_per_s": nb / ms * 1e3 * 1280 * 720 * 128 / 1e9, "parity_ok": ok,
                                           "roofline": {"kernel": "k_bm_fast<15>", "bound": "integer ALU pipe (VABSDIFF4), image-sized HBM traffic",
                                                        "achieved": 4.0 * 1280 * 720 * nb / (ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                                                        "frac": 4.0 * 1280 * 720 * nb / (ms / 1e3) / 1e9 / peak}}
    del bm, L, R
    torch.cuda.empty_cache()
    # the reference's whole compute_disparity (stereo_vision.py:132-184) at the cfg3 settings: left matcher, right-view
    # matcher and the WLS filter.  No reference arm beside it: cv2.ximgproc exists on neither box (WLS parity unpinned).
    settings = dict(window_size=5, min_disparity=0, num_disparities=SGBM_KW["numDisparities"], block_size=SGBM_KW["blockSize"],
                    disp12_max_diff=SGBM_KW["disp12MaxDiff"], uniqueness_ratio=SGBM_KW["uniquenessRatio"],
                    speckle_window_size=SGBM_KW["speckleWindowSize"], speckle_range=SGBM_KW["speckleRange"],
                    pre_filter_cap=SGBM_KW["preFilterCap"], lmbda=80000.0, sigma=1.2, sgbm_mode="HH")
    nb = min(32, int(Ld.shape[0]))
    ms = timed(lambda: sm.compute_disparity(Ld[:nb], Rd[:nb], settings, "SGBM"), 3)
    displ, filt = sm.compute_disparity(Ld[:1], Rd[:1], settings, "SGBM")
    l0, r0 = Ld[0].cpu().numpy(), Rd[0].cpu().numpy()
    ok = bool(np.array_equal(displ[0].cpu().numpy(), cv2.StereoSGBM_create(**dict(SGBM_KW, disp12MaxDiff=1000000, uniquenessRatio=0,
                                                                                  speckleWindowSize=0)).compute(l0, r0)))
    other["compute_disparity_cfg3_left_right_wls"] = {
        "batch": nb, "ms_per_batch": ms, "pairs_per_s": nb / ms * 1e3, "displ_parity_ok": ok,
        "what": "sm.compute_disparity: left SGBM + right-view SGBM + WLS filter, device-resident gray pairs",
        "reference_arm": "none: cv2.ximgproc is not in this image (WLS parity unpinned, DESIGN.md 4.6)"}
    latency = {}
    for nb in (1, 8):
        m = sm.StereoSGBM_create(**SGBM_KW)
        ms = timed(lambda: chain(Ld[:nb], Rd[:nb], m), 20)
        latency["batch_%d" % nb] = {"ms_per_call": ms, "pairs_per_s": nb / ms * 1e3,
                                    "what": "remap x2 + StereoSGBM MODE_HH + speckle + reprojectImageTo3D, device-resident, per call"}
        del m
    return other, latency


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
    stage_sum, sweep_sum = {}, [0.0, 0.0]
    launches0 = lib.b2s_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        chain(Ld, Rd)
        for k, v in matcher.stage_ms().items():       # waits for this step's last stage event
            stage_sum[k] = stage_sum.get(k, 0.0) + v
        s1, s2 = matcher.sweep_ms()
        sweep_sum[0] += s1
        sweep_sum[1] += s2
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
    last = (args.steps - 1) % NS

    # ---- e2e variants, reported BESIDE the dense one (never instead of it): what comes back over the host link ----
    # disp: the int16 disparity only (what matcher.compute returns);  cloud: disparity + the ordered compacted point
    # cloud points_3d[mask] that disparity_calculation.py:308-321 actually consumes (mask = valid disparity)
    e2e_variants = {}
    if not args.no_variants:
        verts_h = torch.empty((B * H * W, 3), dtype=torch.float32).pin_memory()
        vdisp_h = [torch.empty((B, H, W), dtype=torch.int16).pin_memory() for _ in range(NS)]

        def variant_step(i, mode):
            k = i % NS
            with torch.cuda.stream(streams[k]):
                Ls[k].copy_(Lh, non_blocking=True)
                Rs[k].copy_(Rh, non_blocking=True)
                disp, xyz = chain(Ls[k], Rs[k], matchers[k])
                vdisp_h[k].copy_(disp, non_blocking=True)
                nbytes = disp.numel() * 2
                if mode == "cloud":
                    verts, _, offs = sm.point_cloud(disp, xyz, None, lo=0)      # one sync inside: the vertex count
                    verts_h[: verts.shape[0]].copy_(verts, non_blocking=True)
                    nbytes += verts.numel() * 4 + offs.numel() * 8
                keep[k] = (disp, xyz)
            return nbytes

        for mode in ("disp", "cloud"):
            vsteps = max(3, args.steps // 3)
            for i in range(NS):
                variant_step(i, mode)
            e2e_sync()
            barrier()
            e0.record()
            nb = 0
            for i in range(vsteps):
                nb = variant_step(i, mode)
            for st in streams:
                torch.cuda.current_stream().wait_stream(st)
            e1.record()
            barrier()
            ms_v = max_over_ranks(e0.elapsed_time(e1))
            e2e_variants[mode] = {"value": world * B * vsteps / (ms_v / 1e3), "unit": "pairs/s", "steps": vsteps,
                                  "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(nb)}
        del verts_h, vdisp_h

    disp_h, xyz_h = disp_h[last], xyz_h[last]
    # ---- parity gate of the run: EVERY rank checks pairs of the result it just produced against cv2 ------
    parity, parity_detail = None, None
    if not args.no_parity:
        import cv2
        cv2.setNumThreads(max(1, (os.cpu_count() or 1) // world))
        picks = sorted({0, (17 * rank + 5) % B, B - 1})
        ok = True
        for i in picks:
            lr = cv2.remap(Lh[i].numpy(), mx, my, cv2.INTER_LINEAR)
            rr = cv2.remap(Rh[i].numpy(), mx, my, cv2.INTER_LINEAR)
            want = cv2.StereoSGBM_create(**SGBM_KW).compute(lr, rr)
            wxyz = cv2.reprojectImageTo3D(want.astype(np.float32) / 16.0, Q)
            ok = ok and bool(np.array_equal(disp_h[i].numpy(), want) and
                             np.array_equal(xyz_h[i].numpy().view(np.uint32), wxyz.view(np.uint32)))
        flags = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        if world > 1:
            gathered = [torch.zeros_like(flags) for _ in range(world)]
            dist.all_gather(gathered, flags)
            per_rank = [bool(int(t.item())) for t in gathered]
        else:
            per_rank = [ok]
        parity = all(per_rank)
        parity_detail = {"per_rank": per_rank, "pairs_checked_per_rank": len(picks),
                         "what": "int16 disparity equal and float32 XYZ bitwise equal to cv2 on the e2e result"}

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
    roofline = {"bound": "hbm", "kernel": "k_wave8 (2 launches per step: forward sweep of 4 path directions, mirrored sweep of the "
                                          "other 4 with the winner-take-all fused)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes_step / nl,
                "avg_launch_ms": agg_ms / nl, "traffic": traffic, "traffic_source": traffic_src,
                # the two launches apart (an event between them): the forward sweep, and the mirrored one with the fused WTA
                "per_sweep": [{"launch": name, "ms": ms / args.steps, "achieved": alg_bytes_step / nl / (ms / args.steps / 1e3) / 1e9,
                               "frac": alg_bytes_step / nl / (ms / args.steps / 1e3) / 1e9 / peak}
                              for name, ms in (("k_wave8<FINAL=false>: 4 directions, writes S1", sweep_sum[0]),
                                               ("k_wave8<FINAL=true>: 4 directions + S1, winner-take-all", sweep_sum[1])) if ms > 0],
                "stage_ms_per_step": {k: v / args.steps for k, v in stage_sum.items()},
                "stage_share_of_step": {k: v / args.steps / (ms_total / args.steps) for k, v in stage_sum.items()}}

    # ---- free the e2e buffers before the extra measurements ---------------------------------------------------
    wsbytes = matcher.workspace_bytes(H, W, B)
    del keep, Ls, Rs, matchers
    torch.cuda.empty_cache()

    # ---- cfg5 as BASELINE.json states it: --job N DISTINCT seeded pairs, sharded contiguously over the ranks -----
    job = None
    if args.job > 0:
        job = run_job(args, sm, torch, dist, dev, rank, world, chain, matcher, barrier, max_over_ranks, mx, my, Q)

    # ---- optional NCCL gather of the point clouds (north_star: "NCCL is used only to gather point clouds if requested")
    gathered_cloud = None
    if world > 1:
        from stereo_match_b200.batch import gather_host_results
        disp, xyz = chain(Ld[:8], Rd[:8])
        torch.cuda.synchronize()
        e0.record()
        allxyz = gather_host_results(xyz, 8 * world, dst=0)          # CUDA tensors -> ncclAllGather over NVLink
        e1.record()
        torch.cuda.synchronize()
        gms = max_over_ranks(e0.elapsed_time(e1))
        if rank == 0:
            gathered_cloud = {"pairs": 8 * world, "bytes": int(allxyz.numel() * 4), "ms": gms,
                              "GBps": allxyz.numel() * 4 / gms / 1e6, "rank0_block_equal": bool(torch.equal(allxyz[:8], xyz))}
        del allxyz, disp, xyz

    # ---- the other BASELINE configurations and the small-batch ("video frame") latency, rank 0 of a 1-GPU run ------
    other, latency = None, None
    if rank == 0 and world == 1 and not args.no_extras:
        other, latency = run_extras(sm, torch, chain, Ld, Rd, Lh, mx, my, Q, peak)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        # the same shape as `--impl reference --steps 20 --warmup 5`: steps of 2 pairs per worker, a short warm-up
        ksteps = max(1, args.cpu_pairs // 2)
        v, dt, _ = cv2_pairs_per_s(cores, 2, 3, ksteps)
        cpu = {"value": v, "unit": "pairs/s", "cores": cores, "kind": "reference",
               "sample": "%d pairs (%d steps x 2 per persistent worker process, cv2.setNumThreads(1) each, setup and 6 warm-up "
                         "pairs per worker outside the timed region), same chain, %.1f s" % (cores * 2 * ksteps, ksteps, dt)}

    if rank == 0:
        line = {
            "metric": "pairs_per_s", "value": value, "unit": "pairs/s", "gde_per_s": value * GDE_PER_PAIR,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_step_per_gpu": B, "parallelism": "pair-parallel x%d, no collective" % world,
                       "l2": "per-step working set %.1f GB >> 126 MB L2 (no flush needed)" % (wsbytes / 1e9)},
            "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "wall_ms_per_step": wall_e2e / args.steps,
                    "how": "%d streams x 1 matcher handle each" % NS + ", pinned host buffers, H2D + kernels + D2H per step",
                    "variants": e2e_variants},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity_ok": parity, "parity": parity_detail,
            "job": job, "gathered_cloud": gathered_cloud, "other_configs": other, "latency": latency,
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
    ap.add_argument("--ref-pairs", type=int, default=2, help="--impl reference: pairs per worker process and step")
    ap.add_argument("--e2e-streams", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-variants", action="store_true", help="skip the disparity-only / compacted-cloud e2e variants")
    ap.add_argument("--no-extras", action="store_true", help="skip other_configs and the small-batch latency figures")
    ap.add_argument("--job", type=int, default=0, help="cfg5 as BASELINE.json states it: this many DISTINCT pairs, sharded over the ranks")
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
