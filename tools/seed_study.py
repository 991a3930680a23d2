#!/usr/bin/env python
"""Spread of the t = 0 replication value over the initial weights of the network (the only seed besides
the two Sobol engines is the init seed, ours; report >= 5).  C2-shaped runs: 2^m paths x 52 weekly steps, EO net, scaled
mode with the defaults of the bench (8 steps per epoch, 8 virtual shards from 2^18 paths on).

    python tools/seed_study.py [m] > profiles/r02_seed_study.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 20
D = 52
r, sigma, S0, K = 0.05, 0.2, 100.0, 100.0
p = dict(S0=S0, mu=r, sigma=sigma, dt=1.0 / D)
tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True)
bond = orp_b200.bond_curve(r, 1.0, D + 1)
vals = torch.empty(2 ** m, dtype=torch.float32, device="cuda")
sums, _ = orp_b200.payoff_sums(tm, [K], [-1], "call", values_out=vals, value_scale=1 / S0)
rows = []
for seed in (1234, 1, 2, 3, 4, 5, 6, 7):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = orp_b200.fit_replication(tm, bond, vals.clone(), net="eo", scale=S0, init_seed=seed)
    e1.record()
    torch.cuda.synchronize()
    rows.append(dict(init_seed=seed, V0=res["V0"] * S0, phi0=res["phi0"], psi0=res["psi0"],
                     epochs=sum(s["epochs_run"] for s in res["steps"]), first_fit_epochs=res["steps"][0]["epochs_run"],
                     loss_first_fit=res["steps"][0]["loss"], loss_last_fit=res["steps"][-1]["loss"], ms=e0.elapsed_time(e1)))
v = np.array([q["V0"] for q in rows])
print(json.dumps(dict(config=f"2^{m} paths x {D} weekly steps, call K = S0 = 100, r = .05, sigma = .2, EO net, scaled mode",
                      price_mc=float(np.exp(-r) * sums[0] / 2 ** m), V0_mean=float(v.mean()), V0_std=float(v.std(ddof=1)),
                      V0_min=float(v.min()), V0_max=float(v.max()), runs=rows), indent=1))
