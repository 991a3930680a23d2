#!/usr/bin/env python
"""C4-shaped sweep timing: 64 calls (8 strikes x 8 maturities) on one 2^m x 252 GBM set, stored-path route (orp_paths +
orp_payoff_reduce) against the fused generate + payoff kernel (orp_paths_payoff), CUDA events.

    python tools/bench_fused.py [m]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
D = 252
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / D)
strikes = np.repeat(np.linspace(80.0, 120.0, 8), 8)
cols = np.tile(np.arange(1, 9) * 31 + 4, 8)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(fn, reps=10):
    fn()
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


tm = torch.empty((D + 1, 2 ** m), dtype=torch.float64, device="cuda")
ms_paths, _ = timed(lambda: orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, out=tm, time_major=True))
ms_sweep, (ref, _) = timed(lambda: orp_b200.price_european(tm, strikes, cols, "call", r=0.05, dt=1 / D))
for name, ks, ts in (("64 options, 8 maturities", strikes, cols), ("8 options, terminal step", strikes[:8], np.full(8, D)),
                     ("1 option, terminal step", strikes[:1], np.full(1, D))):
    ms, (got, _) = timed(lambda: orp_b200.price_european_fused("gbm_log", p, ks, ts, "call", r=0.05, m=m, steps=D))
    print(f"fused {name:26s} {ms:8.3f} ms  {2 ** m * D / ms / 1e6:8.1f} G path-steps/s  equivalent {2 ** m * D * 8 / ms / 1e6:7.1f} GB/s  "
          f"{len(ks) / ms * 1e3:10.0f} options/s", flush=True)
print(f"stored paths {ms_paths:8.3f} ms ({2 ** m * D / ms_paths / 1e6:.1f} G path-steps/s) + sweep {ms_sweep:.3f} ms = {ms_paths + ms_sweep:.3f} ms", flush=True)
got, _ = orp_b200.price_european_fused("gbm_log", p, strikes, cols, "call", r=0.05, m=m, steps=D)
print("max rel diff fused vs stored:", float(np.abs(got / ref - 1).max()))
