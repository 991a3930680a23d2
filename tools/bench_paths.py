#!/usr/bin/env python
"""Kernel-level timing of the path kernels (CUDA events, inputs far larger than L2): GB/s of kept columns."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
D = int(sys.argv[2]) if len(sys.argv) > 2 else 252
cases = [("gbm_log", "float64", "incremental", 1), ("gbm_log", "float32", "incremental", 1), ("gbm_arith", "float64", "incremental", 1),
         ("sv", "float64", "incremental", 1), ("gbm_log", "float64", "bridge", 1), ("gbm_log", "float64", "incremental", 21)]
P = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / D, s0=0.16679, a=0.0033281299103885727, b=0.1562947229160206, c=0.01583113545983743)
for model, dt, order, keep in cases:
    kw = dict(m=m, steps=D, dtype=dt, ordering=order, keep_every=keep, time_major=True)
    out = orp_b200.simulate_paths(model, P, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        orp_b200.simulate_paths(model, P, out=out, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gb = out.numel() * out.element_size() / 1e9
    print(f"{model:10s} {dt:8s} {order:12s} keep={keep:3d}  {ms:8.3f} ms  {gb / ms * 1e3:8.1f} GB/s  {2 ** m * D / ms / 1e6:8.1f} G path-steps/s", flush=True)
    del out
