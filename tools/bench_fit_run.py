#!/usr/bin/env python
"""Timing of the persistent fit kernel alone (orp_fit_run through ReplicationNet.fit): E epochs of S optimiser steps on
2^m resident samples, with the batch as 1 segment and as 8 segments (the virtual shards of the rank-count-invariant
fit).  Prints ms per epoch, samples/s and the per-step time.

    python tools/bench_fit_run.py [m] [epochs] [steps_per_epoch] [net]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
import orp_b200  # noqa: E402
from orp_b200.fit import ReplicationNet, Optimizer, init_params, scale_column  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 26
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 40
spe = int(sys.argv[3]) if len(sys.argv) > 3 else 8
net = sys.argv[4] if len(sys.argv) > 4 else "eo"
D = 2
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / 252)
tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True)
n = 2 ** m
vals = torch.empty(n, dtype=torch.float32, device="cuda")
orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
x0, x1 = scale_column(tm[D - 1], 0.01), scale_column(tm[D], 0.01)
f1 = torch.ones(n, dtype=torch.float32, device="cuda") if net == "rp" else None
f2 = torch.zeros(n, dtype=torch.float32, device="cuda") if net == "rp" else None
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for cols, label in (((x0, x1, 1.0), "fp32 pre-scaled"), ((tm[D - 1], tm[D], 0.01), "fp64 columns")):
    for segments in (1, 8):
        model = ReplicationNet(net, init_params(net, 1234, 0.1 if net == "eo" else [0.9, 0.1]), "cuda")
        opt = Optimizer("cuda")
        b = model.batch(cols[0], cols[1], vals, cols[2], 1.0, 1.0, feat1=f1, feat2=f2, segments=segments)
        best = 1e30
        for rep in range(3):
            e0.record()
            h = model.fit(b, opt, epochs, 10 ** 6, False, mode="scaled", steps_per_epoch=spe)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        assert h["epochs_run"] == epochs
        print(f"{net} {label:16s} segments={segments}  {best / epochs:8.4f} ms/epoch  {n * epochs / best / 1e6:7.2f} G samples/s  "
              f"{1e3 * best / epochs / spe:7.2f} us/step", flush=True)
        model.close()
