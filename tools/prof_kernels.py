#!/usr/bin/env python
"""Short driver for ncu: a handful of launches of every hot kernel on inputs larger than L2.

    python tools/prof_kernels.py [m] [D]      # default 2^24 paths x 64 steps (fp64 columns: 8.7 GB)
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
D = int(sys.argv[2]) if len(sys.argv) > 2 else 64
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / D)
tm = None
for _ in range(3):
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True, out=tm)
vals = torch.empty(2 ** m, dtype=torch.float32, device="cuda")
for _ in range(3):
    sums, otm = orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
bond = orp_b200.bond_curve(0.05, 1.0, D + 1)
# three time steps of the backward loop, few epochs each
res = orp_b200.fit_replication(tm[-4:], bond[-4:], vals, net="eo", scale=100.0, epochs_first=6, epochs_next=3,
                               total_paths=2 ** m)
torch.cuda.synchronize()
print("ok", float(np.exp(-0.05) * sums[0] / 2 ** m), res["V0"] * 100.0, res["launches"])
