#!/usr/bin/env python
"""BASELINE.json configs C1..C4 end to end on one GPU (C5 is bench.py): device times (CUDA events), prices, errors.

    python tools/run_configs.py > profiles/r02_configs.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402
from oracle import paths as opaths  # noqa: E402  (Black-Scholes closed form only)

dev = torch.device("cuda:0")


def timed(fn, reps=3):
    fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def gbm_case(name, m, D, keep, ordering="incremental", fit=True, spe=8):
    S0, K, T, r, sigma = 100.0, 100.0, 1.0, 0.05, 0.2
    p = dict(S0=S0, mu=r, sigma=sigma, dt=T / D)
    ms_paths, tm = timed(lambda: orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, keep_every=keep, ordering=ordering,
                                                        device=dev, time_major=True))
    vals = torch.empty(2 ** m, dtype=torch.float32, device=dev)
    ms_pay, (sums, otm) = timed(lambda: orp_b200.payoff_sums(tm, [K], [-1], "call", values_out=vals, value_scale=1 / S0))
    mc = float(np.exp(-r * T) * sums[0] / 2 ** m)
    rec = dict(config=name, paths_log2=m, steps=D, kept_columns=tm.shape[0], ordering=ordering, ms_paths=ms_paths,
               path_steps_per_s=2 ** m * D / ms_paths * 1e3, ms_payoff=ms_pay, price_mc=mc,
               black_scholes=opaths.black_scholes_call(S0, K, T, r, sigma))
    if fit:
        bond = orp_b200.bond_curve(r, T, D + 1, keep)
        ms_fit, res = timed(lambda: orp_b200.fit_replication(tm, bond, vals.clone(), net="eo", scale=S0, steps_per_epoch=spe,
                                                             var_quantiles=(.98, .99)), reps=1)
        rec.update(ms_fit=ms_fit, V0_nn=res["V0"] * S0, phi0=res["phi0"], psi0=res["psi0"],
                   epochs=sum(s["epochs_run"] for s in res["steps"]), VaR_t0=res["steps"][-1]["VaR"])
    return rec


out = []
out.append(gbm_case("C1 single time step, 2^16 paths (fine grid 252, rebalancing = T)", 16, 252, 252))
out.append(gbm_case("C2 multi time step, 52 weekly steps, 2^20 paths, incremental", 20, 52, 1))
out.append(gbm_case("C2 multi time step, 52 weekly steps, 2^20 paths, Brownian bridge", 20, 52, 1, ordering="bridge"))

# C3: stochastic volatility (Extra notebook parameters), 252 steps, 2^22 paths, fp64, call K = Y0 = 1 — with the
# notebook's calibrated c and with the c = 0.075 that Replicating_Portfolio_SV itself would read (SURVEY C-4)
for c_sv, tag in ((0.01583113545983743, "c = 0.0158 (notebook calibration)"), (0.075, "c = 0.075 (the module's own parameter, SURVEY C-4)")):
    sv = dict(S0=1.0, mu=0.03, s0=0.16679, a=0.0033281299103885727, b=0.1562947229160206, c=c_sv, dt=1 / 252)
    ms, tm = timed(lambda: orp_b200.simulate_paths("sv", sv, m=22, steps=252, device=dev, time_major=True))
    sums, otm = orp_b200.payoff_sums(tm, [1.0], [-1], "call")
    last = tm[-1]
    out.append(dict(config=f"C3 stochastic-volatility call, 252 steps, 2^22 paths, fp64, {tag}", ms_paths=ms,
                    path_steps_per_s=2 ** 22 * 252 / ms * 1e3, GBs=tm.numel() * 8 / ms / 1e6,
                    price_mc=float(np.exp(-0.03) * sums[0] / 2 ** 22), non_finite_terminal=int((~torch.isfinite(last)).sum()),
                    mean_Y_T=float(torch.nanmean(last))))
    del tm, last

# C4: 64 European options sharing one 2^24-path 252-step GBM Sobol set (one GPU here; shards are independent)
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252)
ms_paths, tm = timed(lambda: orp_b200.simulate_paths("gbm_log", p, m=24, steps=252, device=dev, time_major=True))
strikes = np.repeat(np.linspace(80.0, 120.0, 8), 8)
cols = np.tile(np.arange(1, 9) * 31 + 4, 8)
ms_sweep, (prices, p_otm) = timed(lambda: orp_b200.price_european(tm, strikes, cols, "call", r=0.05, dt=1 / 252))
bs = np.array([opaths.black_scholes_call(100.0, k, c / 252, 0.05, 0.2) for k, c in zip(strikes, cols)])
del tm
# the same sweep from the fused generate + payoff kernel: no path is ever stored
ms_fused, (fprices, _) = timed(lambda: orp_b200.price_european_fused("gbm_log", p, strikes, cols, "call", r=0.05, m=24, steps=252, device=dev),
                               reps=5)
out.append(dict(config="C4 strike/maturity sweep, 64 calls on one 2^24 x 252 GBM set (1 GPU)", ms_paths=ms_paths,
                path_steps_per_s=2 ** 24 * 252 / ms_paths * 1e3, ms_sweep=ms_sweep, options_per_s=64 / ms_sweep * 1e3,
                ms_stored_total=ms_paths + ms_sweep,
                ms_fused_generate_payoff=ms_fused, fused_options_per_s=64 / ms_fused * 1e3,
                fused_equivalent_path_steps_per_s=2 ** 24 * 252 / ms_fused * 1e3, fused_equivalent_GBps=2 ** 24 * 252 * 8 / ms_fused / 1e6,
                fused_max_rel_diff_vs_stored=float(np.abs(fprices / prices - 1).max()),
                max_abs_err_vs_black_scholes=float(np.abs(prices - bs).max()),
                max_rel_err_vs_black_scholes=float(np.abs(prices / bs - 1).max())))
print(json.dumps(out, indent=1))
