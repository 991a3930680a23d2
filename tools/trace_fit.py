#!/usr/bin/env python
"""Phase breakdown of one optimiser step of the persistent fit kernel, on one GPU or — under torchrun — on N GPUs with
the in-kernel peer exchange (2^m paths in total, 8 virtual shards).  Needs the instrumented build:

    python option-replicating-portfolio-with-neural-networks_b200/build.py --trace
    ORP_B200_LIB=$PWD/option-replicating-portfolio-with-neural-networks_b200/liborp_b200_trace.so python tools/trace_fit.py [m] [parts] [segments]
    ORP_B200_LIB=... python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/trace_fit.py 26 8

Blocks 0 and G-1 of every rank write %globaltimer at eight points of each of the first 512 optimiser steps; rank 0
prints the mean time between the points."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402
from orp_b200.fit import ReplicationNet, Optimizer, init_params, scale_column, virtual_segments  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
parts = int(sys.argv[2]) if len(sys.argv) > 2 else 8
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252)
tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=2, shard=(rank, world), device=dev, time_major=True)
vals = torch.empty(tm.shape[1], dtype=torch.float32, device=dev)
orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
mdl = ReplicationNet("eo", init_params("eo", 1, 0.1), dev, peer_exchange=True if world == 1 else None)
opt = Optimizer(dev)
segs = int(sys.argv[3]) if len(sys.argv) > 3 else virtual_segments(tm.shape[1], None, 8)
b = mdl.batch(scale_column(tm[1], 0.01), scale_column(tm[2], 0.01), vals, 1.0, 0.01, 0.01, segments=segs)
for rep in range(2):
    mdl.fit(b, opt, epochs=40, patience=1000, use_schedule=False, steps_per_epoch=parts)
torch.cuda.synchronize()
raw = mdl.scratch.cpu().numpy()[-65536:].view(np.uint64)[:2 * 512 * 8].reshape(2, 512, 8).astype(np.float64)
names = ["accumulate + warp totals (all segments)", "barrier + publish block totals", "lr", "owned columns: poll + sum + publish to ranks",
         "poll the ranks' totals", "(sync)", "adam + callbacks + weight reload"]
n_steps = min(512, 40 * parts)
if rank == 0:
    print(f"world {world}, 2^{m} paths in total, {segs} segment(s) per rank, {parts} optimiser steps per epoch")
    for blk, label in ((0, "block 0"), (1, "block G-1")):
        t = raw[blk, 8:n_steps]                                   # skip the first steps (cold)
        d = np.diff(t, axis=1)
        print(f"{label}: step total {np.mean(t[1:, 0] - t[:-1, 0]) / 1e3:.2f} us")
        for k in range(7):
            print(f"   {names[k]:46s} {d[:, k].mean() / 1e3:7.2f} us")
        print(f"   {'step end -> next start':46s} {np.mean(t[1:, 0] - t[:-1, 7]) / 1e3:7.2f} us")
mdl.close()
if world > 1:
    dist.destroy_process_group()
