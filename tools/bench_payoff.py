#!/usr/bin/env python
"""Kernel-level timing of the payoff-and-reduction kernels (CUDA events): one option on one column, and the C4-shaped
sweep of 64 European calls (8 strikes x 8 maturity columns) off one stored path set.

    python tools/bench_payoff.py [m] [D]        # default 2^24 paths x 64 steps (fp64 columns: 8.7 GB, larger than L2)
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
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for dtype in ("float64", "float32"):
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, dtype=dtype, time_major=True)
    n, esz = 2 ** m, tm.element_size()
    vals = torch.empty(n, dtype=torch.float32, device="cuda")
    strikes = np.repeat(np.linspace(80.0, 120.0, 8), 8)
    cols = np.tile(np.arange(1, 9) * (D // 8), 8)
    cases = (("1 call, terminal column, values written", [100.0], [D], vals, 1, 1),
             ("1 call, terminal column", [100.0], [D], None, 1, 1),
             ("8 calls on one column", np.linspace(80.0, 120.0, 8), [D] * 8, None, 8, 1),
             ("64 calls = 8 strikes x 8 columns", strikes, cols, None, 64, 8))
    for kind in ("call", "floor"):
        for name, K, c, out, n_opt, n_col in cases:
            for rep in range(2):
                e0.record()
                for _ in range(10):
                    sums, otm = orp_b200.payoff_sums(tm, K, c, kind, values_out=out, value_scale=0.01, s0=100.0)
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            gb = n * esz * n_col + (4 * n if out is not None else 0)
            print(f"{dtype} {kind:5s} {name:42s} {ms * 1e3:9.1f} us  {gb / ms / 1e6:8.1f} GB/s  {n_opt / ms * 1e3:10.0f} options/s"
                  f"  sum0 {float(sums[0]) / n:.6f}", flush=True)
    del tm
