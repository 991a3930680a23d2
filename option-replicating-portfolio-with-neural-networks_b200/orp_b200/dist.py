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


def tree_sum_rows(rows):
    """Pairwise (balanced-tree) sum of the rows of a NumPy array in row order: ``((r0 + r1) + (r2 + r3)) + ...`` — the
    association the fit kernels use over virtual shards and ranks (``tree_sum`` in ``csrc/orp_fit.cu``), so a result
    assembled from per-shard sums does not depend on how many ranks held the shards."""
    import numpy as np
    v = [np.asarray(r, dtype=np.float64) for r in rows]
    n, stride = len(v), 1
    while stride < n:
        for i in range(0, n - stride, 2 * stride):
            v[i] = v[i] + v[i + stride]
        stride *= 2
    return v[0]


def allgather_tree_sum_(t, group=None):
    """In-place sum over ranks with the pairwise association of :func:`tree_sum_rows` in rank order (all-gather, then
    every rank adds locally): bit-identical on every rank and, for power-of-two worlds holding aligned groups of
    virtual shards, equal to the one-process sum over all shards."""
    if not is_distributed(group):
        return t
    import torch
    import torch.distributed as dist
    parts = [torch.empty_like(t) for _ in range(dist.get_world_size(group))]
    dist.all_gather(parts, t, group=group)
    n, stride = len(parts), 1
    while stride < n:
        for i in range(0, n - stride, 2 * stride):
            parts[i] = parts[i] + parts[i + stride]
        stride *= 2
    t.copy_(parts[0])
    return t


def allgather_rows(t, group=None):
    """Every rank's copy of ``t`` stacked in rank order, as a NumPy array on the host (``t`` itself for one process)."""
    if not is_distributed(group):
        return t.cpu().numpy()[None]
    import torch
    import torch.distributed as dist
    parts = [torch.empty_like(t) for _ in range(dist.get_world_size(group))]
    dist.all_gather(parts, t, group=group)
    return torch.stack(parts).cpu().numpy()
