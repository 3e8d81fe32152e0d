"""Sharding of a batch of independent sequences over the GPUs of one box.

Sequences never interact (the reference's own parallelism is one process per sequence,
benchmark_results/bench_fft.py:17-21), so the path shards with no data-path collective:
rank r folds its shard on its own GPU, results are gathered on the host in input order.
`torch.distributed` is only used for that host-side gather and for barriers/timing."""
import numpy as np


def estimated_cost(length, max_stack):
    "relative cost of one fold: the beam re-examines O(ms) structures of O(L^2) lag work"
    return float(length) * float(length) * (8.0 + float(max_stack)) + 2000.0


def partition(lengths, world_size, max_stack=50):
    """Greedy longest-processing-time assignment.  Returns a list of index arrays (one per
    rank, ascending) -- deterministic, identical on every rank."""
    lengths = np.asarray(lengths, dtype=np.int64)
    cost = np.array([estimated_cost(l, max_stack) for l in lengths])
    order = np.lexsort((np.arange(len(lengths)), -cost))
    load = np.zeros(world_size)
    shards = [[] for _ in range(world_size)]
    for i in order:
        r = int(np.argmin(load))  # first minimum: deterministic
        shards[r].append(int(i))
        load[r] += cost[i]
    return [np.array(sorted(s), dtype=np.int64) for s in shards]


def plan_ranges(lengths, parts, max_stack=50):
    """The contiguous ranges of the input order that the single-process multi-device entry
    (rafft_fold_batch_multi, `fold_batch(..., devices=[...])`) deals to `parts` devices: parts + 1
    boundaries, equal estimated cost.  Computed by the library (no device needed)."""
    import ctypes as C
    from . import _native as N
    offs = np.zeros(len(lengths) + 1, dtype=np.int64)
    np.cumsum(np.asarray(lengths, dtype=np.int64), out=offs[1:])
    lo = np.zeros(parts + 1, dtype=np.int32)
    rc = N.lib().rafft_plan_ranges(offs.ctypes.data_as(C.POINTER(C.c_int64)), len(lengths), parts, max_stack,
                                   lo.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc != N.RAFFT_OK:
        raise ValueError("rafft_plan_ranges failed: %d" % rc)
    return lo


def imbalance(lengths, shards, max_stack=50):
    "max shard cost / mean shard cost under the cost model"
    c = [sum(estimated_cost(lengths[i], max_stack) for i in s) for s in shards]
    return max(c) / (sum(c) / len(c)) if sum(c) > 0 else 1.0


def fold_batch_sharded(sequences, fold_fn, rank, world_size, max_stack=50, gather=True, group=None):
    """Fold `sequences` (the same full list on every rank) sharded over `world_size` ranks.

    fold_fn(list of sequences) -> list of per-sequence results, for this rank's shard.
    With gather=True rank 0 returns the results of the whole batch in input order (other
    ranks return None); with gather=False every rank returns (indices, its results)."""
    shards = partition([len(s) for s in sequences], world_size, max_stack)
    mine = shards[rank]
    local = fold_fn([sequences[i] for i in mine])
    assert len(local) == len(mine)
    if not gather:
        return mine, local
    if world_size == 1:
        parts = [local]
    else:
        import torch.distributed as dist
        parts = [None] * world_size if rank == 0 else None
        dist.gather_object(local, parts, dst=0, group=group)
        if rank != 0:
            return None
    out = [None] * len(sequences)
    for r, part in enumerate(parts):
        for i, res in zip(shards[r], part):
            out[int(i)] = res
    return out
