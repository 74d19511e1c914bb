"""Point-range sharding of one MSM over the GPUs of a box: one process per GPU under
torch.distributed (NCCL on the GPU box, gloo in the CPU tests).

This mirrors what libff::multi_exp itself does across OpenMP threads (multiexp.tcc:655-687): the
index range is cut into `world` contiguous ranges (the last takes the remainder), every rank
computes the partial sum of its range with the full single-GPU pipeline, the `world` partial group
elements (sizeof(GroupT) bytes each) are gathered, and they are added in rank order
(multiexp.tcc:681-685).  There is no other exchange on the path, so the only collective is one
all_gather of 96..288 bytes per rank.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from . import FORM_SPECIAL, group_bytes, group_sum, multi_exp, shard_range

__all__ = ["world_info", "gather_partials", "combine_partials", "multi_exp_sharded"]


def world_info():
    """(rank, world) of the default process group; (0, 1) when torch.distributed is not in use."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except ImportError:
        pass
    return 0, 1


def gather_partials(partial: np.ndarray) -> np.ndarray:
    """all_gather of one group element per rank -> (world, words) uint64, in rank order."""
    rank, world = world_info()
    partial = np.ascontiguousarray(partial, dtype=np.uint64).reshape(-1)
    if world == 1:
        return partial.reshape(1, -1)
    import torch
    import torch.distributed as dist
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    mine = torch.from_numpy(partial.view(np.int64)).to(dev)
    out = torch.empty(world * mine.numel(), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(out, mine)
    return out.cpu().numpy().view(np.uint64).reshape(world, -1)


def combine_partials(group: int, partials: np.ndarray,
                     sum_fn: Optional[Callable] = None) -> np.ndarray:
    """Sum of the per-rank partial results, in rank order (multiexp.tcc:681-685)."""
    return (sum_fn or group_sum)(group, partials)


def multi_exp_sharded(group: int, bases, scalars, base_form: int = FORM_SPECIAL,
                      partial_fn: Optional[Callable] = None,
                      sum_fn: Optional[Callable] = None) -> np.ndarray:
    """libff::multi_exp over all ranks: `bases` / `scalars` are the FULL host arrays (every rank
    holds them, as a prover process holds its proving key); each rank consumes only its own range.
    Every rank returns the combined result.  `partial_fn(group, bases, scalars, base_form)` and
    `sum_fn(group, points)` default to the device engine; the CPU tests inject checker functions to
    exercise the sharding and the collective without a GPU."""
    rank, world = world_info()
    words = group_bytes(group) // 8
    b = np.ascontiguousarray(bases, dtype=np.uint64).reshape(-1, words)
    n = b.shape[0]
    s = np.ascontiguousarray(scalars, dtype=np.uint64).reshape(n, -1)
    lo, hi = shard_range(n, rank, world)
    part = (partial_fn or multi_exp)(group, b[lo:hi], s[lo:hi], base_form)
    return combine_partials(group, gather_partials(part), sum_fn)
