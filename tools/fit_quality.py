#!/usr/bin/env python
"""How the t=0 replication price depends on the optimiser-steps-per-epoch policy (no reference counterpart:
the reference only ever ran 4096 paths / batch 512 = 8 steps per epoch)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 18
D = int(sys.argv[2]) if len(sys.argv) > 2 else 52
r, sigma, S0, K = 0.05, 0.2, 100.0, 100.0
p = dict(S0=S0, mu=r, sigma=sigma, dt=1.0 / D)
tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True)
bond = orp_b200.bond_curve(r, 1.0, D + 1)
vals = torch.empty(2 ** m, dtype=torch.float32, device="cuda")
sums, _ = orp_b200.payoff_sums(tm, [K], [-1], "call", values_out=vals, value_scale=1 / S0)
mc = float(np.exp(-r) * sums[0] / 2 ** m)
print(f"m={m} D={D} MC price {mc:.4f}")
for spe in (1, 8, 64):
    for seed in (1234, 1, 2):
        t0 = time.time()
        res = orp_b200.fit_replication(tm, bond, vals.clone(), net="eo", scale=S0, steps_per_epoch=spe, init_seed=seed)
        ep = sum(s["epochs_run"] for s in res["steps"])
        print(f"spe={spe:3d} seed={seed:5d} V0={res['V0'] * S0:8.4f} phi0={res['phi0']:.4f} epochs={ep:6d} first={res['steps'][0]['epochs_run']} "
              f"loss_T={res['steps'][0]['loss']:.3e} loss_0={res['steps'][-1]['loss']:.3e} {time.time() - t0:.1f}s", flush=True)
