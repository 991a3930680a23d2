"""world_size-2 gloo runs of the N>1 host logic: contiguous Sobol index shards + sum all-reduce of the
estimator vector reproduce the 1-process result."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import paths as opaths


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path[:0] = [root, os.path.join(root, "option-replicating-portfolio-with-neural-networks_b200")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from orp_b200 import dist as odist, shard_range
    assert odist.is_distributed() and odist.rank_world() == (rank, world)
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 12)
    start, count = shard_range(2 ** 10, rank, world)
    sums, otm = opaths.payoff_sums_window("gbm_log", p, 13, start, count, [90.0, 100.0], [12, 6], "call")
    vec = torch.tensor(np.concatenate([sums, otm, [count]]))
    a = odist.allreduce_sum_(vec.clone())
    b = odist.allgather_ordered_sum(vec.clone())
    # the rank-count-invariant plumbing of the fit (orp_b200/fit.py): this rank holds 8 / world virtual shards; per-shard
    # rows are gathered in rank order and added pairwise, gradient vectors are tree-summed over ranks
    from orp_b200.fit import virtual_segments
    L = virtual_segments(count, None, 8)
    rng = np.random.default_rng(5)
    shard_rows = rng.standard_normal((8, 5)) * 10.0 ** rng.integers(-8, 8, (8, 5))      # wildly different magnitudes
    mine = torch.tensor(shard_rows[rank * L:(rank + 1) * L])
    rows = odist.allgather_rows(mine)
    tree_all = odist.tree_sum_rows(rows.reshape(-1, 5))
    local_tree = torch.tensor(odist.tree_sum_rows(mine.numpy()))
    c = odist.allgather_tree_sum_(local_tree.clone())
    if rank == 0:
        out.put((a.numpy(), b.numpy(), L, tree_all, c.numpy(), shard_rows))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_sums_equal_single_process():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    a, b, L, tree_all, c, shard_rows = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    par = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 12)
    sums, otm = opaths.payoff_sums_window("gbm_log", par, 13, 0, 2 ** 10, [90.0, 100.0], [12, 6], "call")
    np.testing.assert_allclose(a[:2], sums, rtol=1e-13)
    assert np.array_equal(a[2:4], otm) and a[4] == 2 ** 10
    np.testing.assert_array_equal(a, b)
    # 2 ranks x 4 shards: the gathered pairwise sum and "local tree, then tree over ranks" both equal the ONE-process
    # pairwise sum over the 8 shards bit for bit — and differ from the left-to-right sum somewhere in the last bits
    from orp_b200 import dist as odist
    assert L == 4
    one = odist.tree_sum_rows(shard_rows)
    assert np.array_equal(tree_all, one) and np.array_equal(c, one)
    np.testing.assert_allclose(one, shard_rows.sum(axis=0), rtol=1e-12)


def test_single_process_is_a_noop():
    from orp_b200 import dist as odist
    t = torch.ones(3)
    assert not odist.is_distributed() and odist.rank_world() == (0, 1)
    assert odist.allreduce_sum_(t) is t and odist.allgather_ordered_sum(t) is t and odist.allgather_tree_sum_(t) is t
    assert odist.allgather_rows(t).shape == (1, 3)
    rows = np.arange(15.0).reshape(5, 3)                           # odd count: the last row joins at the top of the tree
    assert np.array_equal(odist.tree_sum_rows(rows), ((rows[0] + rows[1]) + (rows[2] + rows[3])) + rows[4])
    from orp_b200.fit import virtual_segments
    assert virtual_segments(2 ** 20, None, 8) == 8 and virtual_segments(2 ** 20 + 4, None, 8) == 1 and virtual_segments(7, None, 8) == 1
