"""Oracle: path stepping, bond curve, column reduction and payoffs (TEST INFRASTRUCTURE).

NumPy restatement of the reference's simulation block.  Every function takes the
``(paths, n_time_steps)`` normal matrix ``W`` the reference calls ``W1`` (column ``t`` drives
step ``t``; column 0 is generated and never used, ``Replicating_Portfolio.py:64``) and works on
an arbitrary window of paths, so sizes the reference cannot materialise (2^22..2^26 paths)
are checked window by window.
"""
from __future__ import annotations

from math import ceil, floor

import numpy as np

from . import sobol


def n_time_steps_of(T: float, dt: float) -> int:
    """``n_time_steps = ceil(T/dt)+1`` (``Replicating_Portfolio.py:51``)."""
    return ceil(T / dt) + 1


def gbm_arith(W: np.ndarray, Y0: float, mu: float, sigma: float, dt: float) -> np.ndarray:
    """Arithmetic Euler, ``Replicating_Portfolio.py:61-65``:
    ``Y_t = Y_{t-1} + Y_{t-1} (mu dt + sigma sqrt(dt) W_t)``."""
    n = W.shape[1]
    Y = np.empty((W.shape[0], n))
    Y[:, 0] = Y0
    for t in range(1, n):
        Y[:, t] = Y[:, t - 1] + Y[:, t - 1] * (mu * dt + sigma * np.sqrt(dt) * W[:, t])
    return Y


def gbm_log(W: np.ndarray, S0: float, r: float, sigma: float, dt: float) -> np.ndarray:
    """Log-Euler + final ``exp``, ``European Options.ipynb`` cell 6 (raw lines 156-165):
    ``lnS_t = lnS_{t-1} + (r - sigma^2/2) dt + sigma sqrt(dt) W_t``."""
    n = W.shape[1]
    S = np.empty((W.shape[0], n))
    S[:, 0] = np.log(S0)
    for t in range(1, n):
        S[:, t] = S[:, t - 1] + ((r - 0.5 * sigma ** 2) * dt + sigma * np.sqrt(dt) * W[:, t])
    return np.exp(S)


def sv_paths(W1: np.ndarray, Wsv: np.ndarray, Y0: float, mu: float, s0: float,
             a: float, b: float, c: float, dt: float, return_vol: bool = False):
    """Stochastic-volatility recursion exactly as ``Replicating_Portfolio.py:277-289``:
    the mean-reversion term has no ``dt``, ``v`` is not truncated and is used as a volatility
    (squared in the drift), the price step uses the NEW ``v_t``; ``exp`` at the end."""
    n = W1.shape[1]
    Y = np.empty((W1.shape[0], n))
    Y[:, 0] = np.log(Y0)
    vt = np.full((W1.shape[0], n), fill_value=float(s0))
    with np.errstate(invalid="ignore"):
        for t in range(1, n):
            vt[:, t] = vt[:, t - 1] + a * (b - vt[:, t - 1]) + c * np.sqrt(vt[:, t - 1] * dt) * Wsv[:, t]
            Y[:, t] = Y[:, t - 1] + ((mu - 0.5 * vt[:, t] ** 2) * dt + vt[:, t] * np.sqrt(dt) * W1[:, t])
    Y = np.exp(Y)
    return (Y, vt) if return_vol else Y


def bond_curve(r: float, T: float, n_time_steps: int) -> np.ndarray:
    """``B = exp(r * linspace(0, T, n_time_steps))`` (``Replicating_Portfolio.py:68``)."""
    return np.exp(r * np.linspace(0, T, n_time_steps))


def reduction_of(n_time_steps: int, T: float, rebalancing: float):
    """Column thinning, ``Replicating_Portfolio.py:92-96``: returns
    ``(reduction, kept_column_indices, reduced_n_time_steps)``."""
    reduction = floor((n_time_steps - 1) / (T / rebalancing))
    cols = np.arange(0, n_time_steps, reduction)
    return reduction, cols, ceil(n_time_steps / reduction)


def payoff(S_T: np.ndarray, K: float, kind: str) -> np.ndarray:
    """CALL / PUT: ``European Options.ipynb`` cell 8 (raw 328-329); FLOOR ``max(Y_T, K)``:
    ``Replicating_Portfolio.py:88``."""
    if kind == "call":
        return np.maximum(S_T - K, 0)
    if kind == "put":
        return np.maximum(K - S_T, 0)
    if kind == "floor":
        return np.where(S_T > K, S_T, K)
    raise ValueError(kind)


def simulate_window(model: str, p: dict, d: int, start: int, count: int,
                    seed: int = 1235, seed_sv: int = 1234, ordering: str = "incremental"):
    """Full-grid paths ``(count, d)`` for an index window, the reference's way: engine of
    dimension ``d = n_time_steps``, normals by ``norm.ppf``, then the step loop."""
    W = sobol.normals_window(d, seed, start, count)
    dt = p["dt"]
    if ordering == "bridge":
        from . import bridge
        W = bridge.bridge_normals(W)
    elif ordering != "incremental":
        raise ValueError(ordering)
    if model == "gbm_arith":
        return gbm_arith(W, p["S0"], p["mu"], p["sigma"], dt)
    if model == "gbm_log":
        return gbm_log(W, p["S0"], p["mu"], p["sigma"], dt)
    if model == "sv":
        Wsv = sobol.normals_window(d, seed_sv, start, count)
        if ordering == "bridge":
            Wsv = bridge.bridge_normals(Wsv)
        return sv_paths(W, Wsv, p["S0"], p["mu"], p["s0"], p["a"], p["b"], p["c"], dt)
    raise ValueError(model)


def payoff_sums_window(model: str, p: dict, d: int, start: int, count: int, strikes, columns,
                       kind: str, **kw):
    """``(sum payoff, count of payoff==0 style OTM)`` per (strike, column) pair over a window.
    OTM follows ``European Options.ipynb`` cell 8 (raw 344, 347): CALL/PUT count ``payoff <= 0``;
    FLOOR follows ``Replicating_Portfolio.py:89``: ``Y_T < Y_0``."""
    S = simulate_window(model, p, d, start, count, **kw)
    sums = np.zeros(len(strikes))
    otm = np.zeros(len(strikes))
    for j, (K, col) in enumerate(zip(strikes, columns)):
        pay = payoff(S[:, col], K, kind)
        sums[j] = pay.sum()
        if kind == "floor":
            otm[j] = np.count_nonzero(S[:, col] < p["S0"])
        else:
            otm[j] = np.count_nonzero(~(pay > 0))
    return sums, otm


def black_scholes_call(S0, K, T, r, sigma):
    from math import erf, exp, log, sqrt
    N = lambda x: 0.5 * (1.0 + erf(x / sqrt(2.0)))
    d1 = (log(S0 / K) + (r + 0.5 * sigma ** 2) * T) / (sigma * sqrt(T))
    return S0 * N(d1) - K * exp(-r * T) * N(d1 - sigma * sqrt(T))
