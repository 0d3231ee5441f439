"""Multi-GPU host logic (one process per GPU, torch.distributed): a neighbourhood's candidates shard
across ranks by contiguous index ranges - every rank holds a replica of the distance table and of
the incumbent path - and the only exchange is the reduction of the chosen move: the smallest global
index of an accepted candidate (the reference's "first improving candidate in enumeration order",
sources/vns_sop_solver.cpp:619,805-806), one int64 all-reduce(MIN).  The result is identical for
any number of ranks because the key is the enumeration index itself.

Backend-agnostic: NCCL on the GPU box, gloo in the CPU tests (tests/test_sharding.py)."""
from __future__ import annotations

import numpy as np

NONE = np.iinfo(np.int64).max


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) of rank's share of n candidates (first n % world ranks get one more)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def global_key(local_first: int, lo: int) -> int:
    """Local first-accept index (-1 = none) -> global enumeration index (NONE = none)."""
    return NONE if local_first < 0 else lo + int(local_first)


def reduce_first_accept(local_first: int, lo: int, group=None, device=None) -> int:
    """all-reduce(MIN) of the global first-accept index; -1 when no rank accepted anything."""
    import torch
    import torch.distributed as dist
    key = torch.tensor([global_key(local_first, lo)], dtype=torch.int64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(key, op=dist.ReduceOp.MIN, group=group)
    k = int(key.item())
    return -1 if k == NONE else k


def reduce_first_accept_(key_tensor, lo: int, group=None):
    """Device-side variant for timed loops: key_tensor is the one-element int64 tensor the engine wrote
    (-1 = none); it is rewritten in place to the global index (NONE = none) and all-reduced without a
    host round trip."""
    import torch
    import torch.distributed as dist
    key_tensor.copy_(torch.where(key_tensor < 0, torch.full_like(key_tensor, NONE), key_tensor + lo))
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(key_tensor, op=dist.ReduceOp.MIN, group=group)
    return key_tensor


def reduce_key_(key_tensor, group=None):
    """The exchange step of vnsgpu_eval_batch_dev_shard: key_tensor is the one-element int64 tensor the engine wrote
    (index_base + local index, or NO_KEY); one all-reduce(MIN), nothing else on the stream."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(key_tensor, op=dist.ReduceOp.MIN, group=group)
    return key_tensor


def sharded_first_accept(eval_fn, cands: np.ndarray, rank: int, world: int, group=None, device=None):
    """Evaluate rank's shard of `cands` with eval_fn(shard) -> (len, lb, flags, local_first) and
    reduce the chosen move.  Returns (lo, hi, len, lb, flags, global_first)."""
    lo, hi = shard_range(len(cands), rank, world)
    ln, lb, fl, first = eval_fn(cands[lo:hi])
    return lo, hi, ln, lb, fl, reduce_first_accept(first, lo, group, device)
