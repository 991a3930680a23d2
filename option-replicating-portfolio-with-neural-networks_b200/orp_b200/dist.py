"""Multi-GPU plumbing: one process per GPU (torchrun), ``torch.distributed`` over NCCL.

The path shards by contiguous Sobol index ranges (:func:`orp_b200.paths.shard_range`): every per-path array
stays on its rank for the whole backward sweep.  The only exchanges are sum-all-reduces of tiny vectors
(estimator sums, and per optimiser step the ``[gradient, loss, count]`` vector of <= 130 doubles), so they
are latency-bound; they are issued on the current stream between the two stages of an epoch.
On CPU (gloo) the same code path is exercised by the world_size-2 tests.
"""
from __future__ import annotations


def is_distributed(group=None) -> bool:
    try:
        import torch.distributed as dist
    except ImportError:      # pragma: no cover
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def rank_world(group=None):
    if not is_distributed(group):
        return 0, 1
    import torch.distributed as dist
    return dist.get_rank(group), dist.get_world_size(group)


def allreduce_sum_(t, group=None):
    """In-place sum over ranks (no-op for a single process)."""
    if is_distributed(group):
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def allgather_ordered_sum(t, group=None):
    """Rank-count-invariant alternative: gather every rank's partial and add them in rank order, so all
    ranks hold bit-identical sums whatever algorithm the collective library would pick for a reduce."""
    if not is_distributed(group):
        return t
    import torch
    import torch.distributed as dist
    parts = [torch.empty_like(t) for _ in range(dist.get_world_size(group))]
    dist.all_gather(parts, t, group=group)
    acc = parts[0].clone()
    for p in parts[1:]:
        acc += p
    t.copy_(acc)
    return t
