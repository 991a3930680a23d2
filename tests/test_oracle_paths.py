"""The oracle's path / payoff restatement against the reference's own outputs (fixtures made by running the
unmodified reference, tests/golden/make_golden.py) and the numbers printed in its notebooks."""
import json
import os

import numpy as np
import pytest

from oracle import bridge, paths, sobol


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_rp_paths_match_reference(golden_dir):
    g = _load(golden_dir, "rp_paths.npz")
    p = json.loads(str(g["params"]))
    d = paths.n_time_steps_of(p["T"], p["dt"])
    sim = dict(S0=p["Y"], mu=p["mu"], sigma=p["sigma"], dt=p["dt"])
    Y = paths.simulate_window("gbm_arith", sim, d, 0, 2 ** p["n_paths"])
    red, cols, n_red = paths.reduction_of(d, p["T"], p["rebalancing"])
    assert red == int(g["reduction"]) and n_red == int(g["n_time_steps"])
    np.testing.assert_array_equal(Y[:, cols], g["Y_reduced"])          # same NumPy ops -> bit-exact
    np.testing.assert_array_equal(Y[:, ::125], g["Y_full_cols"])
    np.testing.assert_array_equal(paths.bond_curve(p["r"], p["T"], d)[cols], g["B_reduced"])
    np.testing.assert_array_equal(paths.payoff(Y[:, -1], p["K"], "floor"), g["Payoff_Y"])
    assert np.mean(Y[:, -1] < p["Y"]) == float(g["out_of_money_P"])


@pytest.mark.parametrize("tag,c", [("c075", 0.075), ("c0158", 0.01583113545983743)])
def test_sv_paths_match_reference(golden_dir, tag, c):
    g = _load(golden_dir, "rp_sv_paths.npz")
    p = json.loads(str(g["params"]))
    d = paths.n_time_steps_of(p["T"], p["dt"])
    W1 = sobol.normals_window(d, 1235, 0, 2 ** p["n_paths"])
    Wsv = sobol.normals_window(d, 1234, 0, 2 ** p["n_paths"])
    Y, vt = paths.sv_paths(W1, Wsv, p["Y"], p["mu"], p["s0"], p["a"], p["b"], c, p["dt"], return_vol=True)
    np.testing.assert_array_equal(Y[:, ::125], g[f"{tag}_Y_full_cols"])
    np.testing.assert_array_equal(vt[:, ::125], g[f"{tag}_vt_cols"])


def test_eo_paths_match_reference(golden_dir):
    g = _load(golden_dir, "eo_paths.npz")
    d = 366
    sim = dict(S0=float(g["S0"]), mu=float(g["r"]), sigma=float(g["sigma"]), dt=float(g["dt_fine"]))
    S = paths.simulate_window("gbm_log", sim, d, 0, 2 ** int(g["n_paths_log2"]))
    red, cols, n_red = paths.reduction_of(d, float(g["T"]), 1 / 52)
    assert red == int(g["reduction"]) and n_red == int(g["n_time_steps"])
    np.testing.assert_array_equal(S[:, cols], g["S_weekly"])
    np.testing.assert_array_equal(S[:, ::73], g["S_fine_cols"])
    np.testing.assert_array_equal(paths.payoff(S[:, cols][:, -1], float(g["K"]), "call"), g["payoff"])


def test_notebook_printed_numbers(golden_dir):
    """KATs from the saved notebook outputs; 5e-7 relative because the author's SciPy/NumPy build differs
    from the installed one in the last printed digits."""
    s = json.load(open(os.path.join(golden_dir, "summary.json")))
    nb = s["notebook_printed"]
    # Single Time Step.ipynb: 8192 paths, dt = 1/12, T = 10, mu = .08, sigma = .15, arithmetic Euler
    Y = paths.simulate_window("gbm_arith", dict(S0=1.0, mu=0.08, sigma=0.15, dt=1 / 12), 121, 0, 8192)
    assert abs(Y[:, -1].mean() / nb["STS_mean_Y_T"] - 1) < 5e-7
    # European Options.ipynb: 4096 x 366 log-Euler
    S = paths.simulate_window("gbm_log", dict(S0=100.0, mu=0.08, sigma=0.15, dt=1 / 365), 366, 0, 4096)
    assert abs(S[:, -1].mean() / nb["EO_mean_S_T_fine"] - 1) < 5e-7
    _, cols, _ = paths.reduction_of(366, 1.0, 1 / 52)
    pay = paths.payoff(S[:, cols][:, -1], 100.0, "call")
    assert abs(S[:, cols][:, -1].mean() / nb["EO_mean_S_T_weekly"] - 1) < 5e-7
    assert abs(pay.mean() / nb["EO_payoff_mean"] - 1) < 2e-6
    assert abs(pay.max() / nb["EO_payoff_max"] - 1) < 5e-7
    assert abs(np.mean(~(pay > 0)) - nb["EO_prob_otm"]) < 1e-4
    # and against the reference run in this container
    assert S[:, -1].mean() == s["eo_mean_S_T_fine"] and pay.mean() == s["eo_payoff_mean"]


def test_black_scholes_kat():
    assert abs(paths.black_scholes_call(100, 100, 1, .05, .2) - 10.450583572185565) < 1e-12


def test_bridge_schedule_and_covariance():
    for D in (1, 2, 3, 7, 52, 252):
        node, left, right, dim = bridge.bridge_schedule(D)
        assert sorted(node.tolist()) == list(range(1, D + 1)) and dim.tolist() == list(range(1, D + 1))
    # Cov(W_s, W_t) = min(s, t): push iid normals through the bridge map
    rng = np.random.default_rng(0)
    D = 16
    Z = bridge.bridge_normals(rng.standard_normal((200_000, D + 1)))
    W = np.cumsum(Z[:, 1:], axis=1)
    cov = W.T @ W / W.shape[0]
    expect = np.minimum.outer(np.arange(1, D + 1), np.arange(1, D + 1))
    assert np.abs(cov - expect).max() < 0.15


def test_bridge_price_converges_to_black_scholes():
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 52)
    bs = paths.black_scholes_call(100, 100, 1, .05, .2)
    for ordering in ("incremental", "bridge"):
        S = paths.simulate_window("gbm_log", p, 53, 0, 2 ** 14, ordering=ordering)
        price = np.exp(-0.05) * paths.payoff(S[:, -1], 100.0, "call").mean()
        assert abs(price - bs) < 0.02, (ordering, price, bs)


def test_minus_inf_propagation():
    # window holding the exact-zero coordinate
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252)
    S = paths.simulate_window("gbm_log", p, 253, 2_631_230, 2)
    assert S[0, 235] > 0 and np.all(S[0, 236:] == 0.0) and np.all(np.isfinite(S[1]))
    Y = paths.simulate_window("gbm_arith", p, 253, 2_631_230, 1)
    assert Y[0, 236] == -np.inf and np.all(~np.isfinite(Y[0, 236:]))


@pytest.mark.parametrize("tag", ["rp", "rp_sv", "0.05", "0.1", "0.15", "0.2", "0.3"])
def test_step_lines_of_the_saved_notebook_run(golden_dir, tag):
    """``Multi Time Step.ipynb`` cells 28 / 32 saved what ``RP.Replicating_Portfolio(_SV)`` printed at every rebalancing
    step: ``>> Y_(t) = mean of the fund column, N_(t) = ...`` (RP.py:194).  The oracle's paths on the notebook's parameters
    must give the 40 fund means digit for digit (three decimals as printed; the SV cell's dict literal has two 'c' keys,
    the last one, 0.075, is what the module saw)."""
    import re
    with open(os.path.join(golden_dir, "summary.json")) as f:
        log = json.load(f)["notebook_rp_log"]
    log = log[tag] if tag in log else log["sigma_sweep"][tag]    # cell 29: the same call for five volatilities
    d = paths.n_time_steps_of(10, 1 / 100)
    if tag != "rp_sv":
        sigma = 0.15965 if tag == "rp" else float(tag)
        Y = paths.simulate_window("gbm_arith", dict(S0=1.0, mu=0.09464, sigma=sigma, dt=1 / 100), d, 0, 4096)
    else:
        Y = paths.simulate_window("sv", dict(S0=1.0, mu=0.09620, s0=0.16679, a=0.0033281299103885727,
                                             b=0.1562947229160206, c=0.075, dt=1 / 100), d, 0, 4096)
    red, cols, n_red = paths.reduction_of(d, 10, 1 / 4)
    assert red == 25 and n_red == 41
    means = Y[:, cols].mean(axis=0)
    printed = [re.match(r">> Y_\(([0-9.]+)\) = ([0-9.]+),", line).groups() for line in log["step_lines"]]
    assert len(printed) == 40
    for k, (t, y) in enumerate(printed):                          # the loop runs backwards from t = T
        t_i = n_red - 2 - k
        assert f"{(t_i + 1) * 0.25:.2f}" == t
        assert f"{means[t_i + 1]:.3f}" == y, (t, means[t_i + 1], y)
