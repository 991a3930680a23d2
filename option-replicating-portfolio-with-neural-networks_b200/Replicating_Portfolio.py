"""Drop-in for the reference's ``Replicating_Portfolio.py``: same two entry points, same ``params`` dict,
same return value ``(phi, psi)`` at t = 0 scaled by ``N * P`` — computed on a B200.

    import Replicating_Portfolio as RP ; reload(RP)
    phi, psi = RP.Replicating_Portfolio(parameters)          # Multi Time Step.ipynb cell 28
    phi, psi = RP.Replicating_Portfolio_SV(sv_parameters)    # cell 32

What runs where (reference lines in brackets):

* fund paths ``Y`` [:60-65 / :274-289], mortality intensity ``lambda`` [:72-76], payoff and summary
  statistics [:87-89], the per-step MLP fits, roll-back and hedge-weight means [:193-235]: CUDA kernels of
  ``liborp_b200.so``;
* survivors ``N_t`` [:78-84]: up to 2^16 paths NumPy's legacy ``RandomState.binomial`` on the host, re-seeded per
  step as in the reference — a sequential variable-rejection stream that only NumPy can reproduce bit for bit;
  beyond that (or with ``survivors='device'``) the same binomial thinning from a counter-based generator on the GPU;
* the per-step ``VaR`` quantiles [:120-122]: radix select on the device (``orp_b200.quantiles``);
* plots, DataFrames, per-path Python lists [:117-123, 229-231]: not reproduced; the means they end in are
  reduced on the device.

Imports neither TensorFlow nor matplotlib.  Optional keys beyond the reference's: ``mode`` (``'scaled'``
default, ``'keras'`` = serial batch-512 mini-batches), ``init_seed``, ``device``, ``verbose``, ``survivors``,
``nonfinite`` (``'propagate'`` = the reference's behaviour, default; ``'skip'`` leaves paths that turned NaN / inf —
exact-zero Sobol coordinates under the arithmetic Euler step, from 2^22 paths on — out of the fits).
Like the reference, importing the module seeds NumPy's global generator with 1234 [:27].
"""
from __future__ import annotations

from math import ceil, floor

import numpy as np
from numpy.random import seed

seed(1234)

COST_OF_CAPITAL = .1


def _simulate_common(params, model, sim_params, n_time_steps, reduction, device):
    """Fund paths at the kept columns plus the true terminal column (``Payoff_Y`` is taken before the column
    slice, ``:88`` vs ``:93``)."""
    from orp_b200 import simulate_paths
    m, steps = params['n_paths'], n_time_steps - 1
    kept = simulate_paths(model, sim_params, m=m, steps=steps, seed=1235, keep_every=reduction,
                          device=device, time_major=True)
    if steps % reduction == 0:
        terminal = kept[-1]
    else:
        terminal = simulate_paths(model, sim_params, m=m, steps=steps, seed=1235, keep_every=steps,
                                  device=device, time_major=True)[-1]
    return kept, terminal


def _mortality(params, n_time_steps, dt, reduction, device):
    """``L_paths`` on the device (seed 1234, ``:72-76``) and ``N_paths / N`` at the kept columns (``:78-84, 93``).

    ``params['survivors']``: ``'host'`` — NumPy's legacy ``RandomState.binomial`` re-seeded per step exactly as the
    reference does (bit-identical ``N_paths``, but a host loop over a ``(steps, paths)`` array); ``'device'`` — the
    same binomial thinning drawn from a counter-based generator on the GPU (distributional parity, any path count);
    default: host up to 2^16 paths, device beyond."""
    import torch
    from orp_b200 import simulate_paths, survivors
    m = params['n_paths']
    L_tm = simulate_paths('linear', {'S0': params['l0'], 'mu': params['c'], 'sigma': params['ita'], 'dt': dt},
                          m=m, steps=n_time_steps - 1, seed=1234, device=device, time_major=True)
    where = params.get('survivors', 'host' if m <= 16 else 'device')
    if where == 'device':
        return L_tm, survivors(L_tm, params['N'], dt, keep_every=reduction, seed=1234).to(torch.float64)
    if where != 'host':
        raise ValueError("params['survivors'] must be 'host' or 'device'")
    L_host = L_tm.cpu().numpy()                                   # (n_time_steps, paths)
    N_paths = np.empty((n_time_steps, 2 ** m), dtype=int)
    N_paths[0] = params['N']
    for t in range(1, n_time_steps):
        probabilities = np.exp(-L_host[t] * dt)
        np.random.seed(1234 + t)                                  # as the reference: reproducible draws
        N_paths[t] = np.random.binomial(N_paths[t - 1], probabilities)
    return L_tm, (torch.from_numpy(N_paths[::reduction]).to(torch.float64) / params['N']).to(device)


def _run(params, model, sim_params):
    import torch
    from orp_b200 import fit_replication, payoff_sums, bond_curve

    Y, K, T, r = params['Y'], params['K'], params['T'], params['r']
    N, P = params['N'], params['P']
    rebalancing = params['rebalancing']
    dt = params['dt']
    device = torch.device(params.get('device', 'cuda'))
    ADJUSTMENT_FACTOR = N * P

    n_time_steps = ceil(T / dt) + 1
    reduction = floor((n_time_steps - 1) / (T / rebalancing)) ; print(f'reduction = {reduction}')

    Y_tm, Y_T = _simulate_common(params, model, sim_params, n_time_steps, reduction, device)
    L_tm, N_NN = _mortality(params, n_time_steps, dt, reduction, device)      # N_paths / N at the kept columns
    B = bond_curve(r, T, n_time_steps, reduction)
    n_red = Y_tm.shape[0]
    dt_red = dt * reduction

    # summary statistics and terminal value  [:87-89, :182-185]
    sums, otm = payoff_sums(Y_T.unsqueeze(0).contiguous(), [K], [0], kind='floor', s0=Y)
    n_paths = Y_tm.shape[1]
    out_of_money_P = float(otm[0]) / n_paths
    Payoff_Y = torch.where(Y_T > K, Y_T, torch.full_like(Y_T, float(K)))
    terminal = (Payoff_Y * N_NN[-1]).to(torch.float32)

    N_feat = N_NN.to(torch.float32).contiguous()
    L_feat = L_tm[:n_red].to(torch.float32).contiguous()           # fine-grid column t_i, as the reference indexes it [:200]

    def features(t_i):
        return N_feat[t_i], L_feat[t_i]

    def progress(rec):
        t_i = rec['t_i']
        print(f'\n>> Y_({(t_i + 1) * dt_red:.2f}) = {float(Y_tm[t_i + 1].mean()):.3f}, '
              f'N_({(t_i + 1) * dt_red:.2f}) = {float(N_NN[t_i + 1].mean()):.3f}')

    def risk(rec):                                                 # the reference's per-step print [:122]
        print(f'VaR: {rec["VaR"][0]:4f} (98%),  {rec["VaR"][1]:4f} (99%)')

    chatty = params.get('verbose', True)
    res = fit_replication(Y_tm, B, terminal, net='rp', mode=params.get('mode', 'scaled'), scale=1.0,
                          out_bias=[1 - out_of_money_P, out_of_money_P], init_seed=params.get('init_seed', 1234),
                          features=features, second_model=True, cost_of_capital=COST_OF_CAPITAL,
                          verbose=False, on_step=progress if chatty else None,
                          var_quantiles=(.98, .99) if chatty else None, on_step_end=risk if chatty else None,
                          skip_nonfinite=params.get('nonfinite', 'propagate') == 'skip')
    phi = np.float64(res['phi0'] * ADJUSTMENT_FACTOR)
    psi = np.float64(res['psi0'] * ADJUSTMENT_FACTOR)
    return phi, psi


def Replicating_Portfolio(params):
    """Constant-volatility fund, arithmetic Euler [``Replicating_Portfolio.py:29-235``]."""
    sim = {'S0': params['Y'], 'mu': params['mu'], 'sigma': params['sigma'], 'dt': params['dt']}
    return _run(params, 'gbm_arith', sim)


def Replicating_Portfolio_SV(params):
    """Stochastic-volatility fund [``Replicating_Portfolio.py:237-459``].  As in the reference, ``params['c']``
    is read both as the volatility-of-volatility and as the mortality drift (the notebook's dict has two
    ``'c'`` keys and the last one wins); pass ``c_sv`` to give the volatility process its own value."""
    sim = {'S0': params['Y'], 'mu': params['mu'], 's0': params['s0'], 'a': params['a'], 'b': params['b'],
           'c': params.get('c_sv', params['c']), 'dt': params['dt']}
    return _run(params, 'sv', sim)
