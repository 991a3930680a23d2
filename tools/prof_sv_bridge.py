#!/usr/bin/env python
"""Short driver for ncu: two launches each of the stochastic-volatility and the Brownian-bridge path kernels on C3's
shape (2^22 paths x 252 steps, fp64 columns: 8.5 GB, far larger than L2).

    python tools/prof_sv_bridge.py [m] [D]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
D = int(sys.argv[2]) if len(sys.argv) > 2 else 252
P = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / D, s0=0.16679, a=0.0033281299103885727, b=0.1562947229160206, c=0.01583113545983743)
for model, order in (("sv", "incremental"), ("gbm_log", "bridge")):
    kw = dict(m=m, steps=D, dtype="float64", ordering=order, keep_every=1, time_major=True)
    out = orp_b200.simulate_paths(model, P, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    orp_b200.simulate_paths(model, P, out=out, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"{model:8s} {order:12s} {ms:8.3f} ms  {2 ** m * D / ms / 1e6:8.1f} G path-steps/s  mean terminal {float(out[-1].mean()):.6f}", flush=True)
    del out
