"""Sharding of the (frequency, candidate) batch over the GPUs of a box.

The systems of a sweep are independent (`jax.vmap` over omega, `Problem.py:516`; OpenMP over
batch items, `InnerState.h:277`), so every rank owns a contiguous block of the frequency
list, holds a replica of the pattern / components, and the only exchange is ONE all-reduce
of [loss, grad[P]] (or of the FRF vector for `solveForward`).  torch.distributed is the
plumbing: NCCL over NVLink on the GPUs, gloo in the CPU tests."""
from __future__ import annotations

import numpy as np


def _dist():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist
    except ImportError:
        pass
    return None


def shard_range(n: int, rank: int = None, world: int = None):
    """Contiguous block [lo, hi) of range(n) owned by `rank`; blocks differ by at most one."""
    dist = _dist()
    if rank is None:
        rank = dist.get_rank() if dist else 0
    if world is None:
        world = dist.get_world_size() if dist else 1
    return (n * rank) // world, (n * (rank + 1)) // world


def allreduce_sum(vec: np.ndarray, cuda_device: int = None) -> np.ndarray:
    """Sum a small host vector over all ranks (identity without a process group)."""
    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return vec
    import torch
    t = torch.from_numpy(np.ascontiguousarray(vec))
    if dist.get_backend() == "nccl":
        t = t.to(torch.device("cuda", cuda_device if cuda_device is not None else torch.cuda.current_device()))
    dist.all_reduce(t)
    return t.cpu().numpy()


def sharded_value_and_grad(local_eval, n_items: int, n_params: int, cuda_device: int = None):
    """`local_eval(lo, hi) -> (loss_partial, grad_partial[P])` on this rank's block, summed
    over the ranks with a single all-reduce of 1 + P doubles."""
    lo, hi = shard_range(n_items)
    acc = np.zeros(1 + n_params)
    if hi > lo:
        val, grad = local_eval(lo, hi)
        acc[0] = val
        acc[1:] = grad
    acc = allreduce_sum(acc, cuda_device)
    return float(acc[0]), acc[1:]


def sharded_gather(local_eval, n_items: int, cuda_device: int = None, dtype=np.float64):
    """`local_eval(lo, hi) -> values[hi - lo]`; every rank receives the full vector."""
    lo, hi = shard_range(n_items)
    full = np.zeros(n_items, dtype=dtype)
    if hi > lo:
        full[lo:hi] = local_eval(lo, hi)
    return allreduce_sum(full, cuda_device)
