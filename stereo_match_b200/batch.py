"""Pair-level data parallelism: every stereo pair is independent, so a job of N
pairs is split into contiguous blocks, one per rank (one process per GPU).  There is no
collective on the data path; `gather_*` exists only for callers that want every rank's
results on one rank (north_star: "NCCL is used only to gather point clouds if requested")."""
import torch
import torch.distributed as dist


def shard_range(n_pairs, rank, world):
    """[lo, hi) of the contiguous block of pairs owned by `rank`; sizes differ by at most 1."""
    base, rem = divmod(int(n_pairs), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_host_results(local, n_pairs, dst=0):
    """Gathers per-rank result blocks (leading axis = local pairs, CPU or CUDA tensors) in global
    pair order on rank `dst`.  Uses the process group's backend: NCCL for CUDA tensors (over
    NVLink), gloo for CPU tensors.  Returns the full tensor on `dst`, None elsewhere."""
    rank, world = dist.get_rank(), dist.get_world_size()
    sizes = [shard_range(n_pairs, r, world) for r in range(world)]
    maxn = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((maxn,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    dtype, shape = pad.dtype, pad.shape
    pad = pad.reshape(-1).view(torch.uint8)          # bytes: gloo has no int16
    if dist.get_backend() == "nccl":
        out = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(out, pad)
    else:
        out = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
        dist.gather(pad, out, dst=dst)
    if rank != dst:
        return None
    out = [o.view(dtype).reshape(shape) for o in out]
    return torch.cat([o[: hi - lo] for o, (lo, hi) in zip(out, sizes)], dim=0)
