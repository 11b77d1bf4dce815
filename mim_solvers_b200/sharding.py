"""Multi-GPU plumbing of the batched QP engine.

The batch shards trivially: B independent problems, no exchange step inside a solve.  One
process per GPU (torchrun) owns a contiguous block of ceil(B/G) problems with its own
`mimgpu_ctx`; the only collectives run AFTER the direction kernels, on
`torch.distributed` (NCCL over NVLink on the GPU box, gloo in the CPU tests):
  1. all-reduce of the 5-number batch summary (`mimgpu_summary`, include/mimgpu.h): max KKT,
     sum / max qp_iters, #converged, #failed -- the batch-wide convergence report and the
     "does any shard need a regularisation retry" decision of a sharded solve();
  2. optional gather of dx / du / lag_mul (y, z) to one rank when the caller wants one
     contiguous result.
"""
from __future__ import annotations

import numpy as np


def shard_range(B: int, rank: int, world: int) -> tuple[int, int]:
    """[start, stop) of the problems rank `rank` owns: contiguous blocks of ceil(B/world)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    per = -(-B // world)
    return min(B, rank * per), min(B, (rank + 1) * per)


def _dist():
    import torch.distributed as dist
    return dist


def allreduce_summary(local: dict, device=None, group=None) -> dict:
    """Batch-wide summary from the per-shard `BatchEngine.summary()` dicts.  Two tiny
    collectives (MAX over [max_kkt, max_qp_iters], SUM over the three counters); identity
    when torch.distributed is not initialised (single GPU)."""
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return dict(local)
    import torch
    mx = torch.tensor([local["max_kkt"], float(local["max_qp_iters"])], dtype=torch.float64, device=device)
    sm = torch.tensor([local["sum_qp_iters"], local["n_converged"], local["n_failed"]], dtype=torch.int64,
                      device=device)
    # NaN-propagating max like the device reduction: NCCL/gloo MAX drop NaNs, so count them
    nan = torch.tensor([1 if np.isnan(local["max_kkt"]) else 0], dtype=torch.int64, device=device)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(sm, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(nan, op=dist.ReduceOp.SUM, group=group)
    return dict(max_kkt=float("nan") if int(nan.item()) else float(mx[0].item()),
                max_qp_iters=int(mx[1].item()), sum_qp_iters=int(sm[0].item()),
                n_converged=int(sm[1].item()), n_failed=int(sm[2].item()))


def allreduce_max(value: float, device=None, group=None) -> float:
    """Max over ranks of a device-measured time (multi-GPU numbers are max over ranks)."""
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    import torch
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def gather_field(local, B_total: int, dst: int = 0, group=None):
    """Gathers one solver member ([B_local, ...] torch tensor, CPU or CUDA) of every shard to
    rank `dst` in problem order; returns the [B_total, ...] tensor there and None elsewhere.
    Shards are the contiguous blocks of `shard_range` (the last ones may be shorter / empty)."""
    import torch
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    per = -(-B_total // world)
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    # all_gather of equal-sized blocks works on both NCCL and gloo; dst keeps the result
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    if rank != dst:
        return None
    return torch.cat(parts, dim=0)[:B_total]


__all__ = ["shard_range", "allreduce_summary", "allreduce_max", "gather_field"]
