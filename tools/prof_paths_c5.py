#!/usr/bin/env python
"""Two launches of the C5 path kernel (2^m paths x D steps, fp64, all columns) and two of the fused generate + payoff
kernel on the same index range (64 calls, 8 maturities: no path is stored), and nothing else — for ncu."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 26
D = int(sys.argv[2]) if len(sys.argv) > 2 else 252
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / D)
out = None
for _ in range(2):
    out = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True, out=out)
strikes = np.repeat(np.linspace(80.0, 120.0, 8), 8)
steps = np.tile(np.arange(1, 9) * (D // 8) + D % 8, 8)
for _ in range(2):
    sums, otm = orp_b200.payoff_sums_fused("gbm_log", p, strikes, steps, "call", m=m, steps=D)
torch.cuda.synchronize()
print("ok", float(out[-1, :1024].mean()), float(sums[0]) / 2 ** m)
