#!/usr/bin/env python
"""The reference's own entry point at a size the reference cannot reach: RP.Replicating_Portfolio(parameters) with the
notebook's dictionary (Multi Time Step.ipynb cell 28) but 2^m paths instead of 2^12."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
import Replicating_Portfolio as RP  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
P = {'Y': 1, 'K': 1, 'T': 10, 'mu': 0.09464, 'r': 0.03, 'sigma': 0.15965, 'rebalancing': 1 / 4,
     'N': 10_000, 'P': 100, 'x': 55.0, 'l0': 0.01, 'c': 0.075, 'ita': 0.000597, 'dt': 1 / 100, 'n_paths': m, 'verbose': False}
for fn, ref, extra in ((RP.Replicating_Portfolio, (634_349, 350_176), {}),
                       (RP.Replicating_Portfolio, (634_349, 350_176), {'nonfinite': 'skip'})):
    for rep in range(2):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        phi, psi = fn(dict(P, **extra))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(f"{fn.__name__} {extra} 2^{m} paths x 1000 fine steps, 40 rebalancing steps x 2 fits: phi0 = {phi:,.0f} psi0 = {psi:,.0f} "
          f"(notebook at 2^12 paths: {ref[0]:,} / {ref[1]:,})  {dt:.2f} s", flush=True)
