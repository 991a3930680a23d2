"""The fit oracle (oracle/fit.py, torch-CPU restatement of the Keras semantics) against the ONLY evidence of the
reference's own fit that exists: the training log and prints that its TensorFlow run left in the saved
``European Options.ipynb`` (cell 13: 52 ``model1.fit`` calls on 4096 paths, batch 512; cells 18, 20), extracted by
``tests/golden/make_golden.py::notebook_fit_log`` into ``summary.json``.

TensorFlow's initialiser and shuffle streams cannot be reproduced, so the comparison is statistical and the tolerances
are the method's own seed-to-seed spread (init seeds 1, 2, 3 give V0 = 12.2 .. 13.1 where 1234 gives 11.8):
losses 10 %, epoch counts 30 %, t = 0 price 15 %, hedge ratio 0.02."""
import json
import os

import numpy as np
import pytest

from oracle import fit as ofit, paths as opaths


@pytest.fixture(scope="module")
def eo_run(golden_dir):
    with open(os.path.join(golden_dir, "summary.json")) as f:
        log = json.load(f)["notebook_fit_log"]
    z = np.load(os.path.join(golden_dir, "eo_paths.npz"))
    S0, r, sigma, T = float(z["S0"]), float(z["r"]), float(z["sigma"]), float(z["T"])
    d = 366                                                     # ceil(T / dt) + 1, EO cell 6
    S = opaths.simulate_window("gbm_log", dict(S0=S0, mu=r, sigma=sigma, dt=float(z["dt_fine"])), d, 0, log["paths"])
    red, cols, n_red = opaths.reduction_of(d, T, 1 / 52)
    assert red == int(z["reduction"]) and n_red == log["columns"]
    S = np.ascontiguousarray(S[:, cols])
    B = np.exp(r * np.linspace(0, T, d))[cols]
    # the arrays the notebook's Keras saw (the fixture keeps the rows of a 2^7-path run of the same cells: the first
    # 128 points of the same scrambled Sobol set)
    k = z["S_weekly"].shape[0]
    np.testing.assert_array_equal(S[:k], z["S_weekly"])
    np.testing.assert_allclose(B, z["B_weekly"], rtol=1e-15)
    pay = opaths.payoff(S[:, -1], float(z["K"]), "call")
    import torch
    torch.set_num_threads(1)
    res = ofit.backward_replication(S, B, pay / S0, net="eo", scale=S0, init_seed=1234, shuffle_seed=1234)
    return log, res, S0, float(np.exp(-r * T) * pay.mean())


def _last_loss(h):
    return h["loss"] if "loss" in h else h["losses"][-1]


def test_discounted_payoff_is_the_notebooks(eo_run, golden_dir):
    log, _, _, disc = eo_run
    with open(os.path.join(golden_dir, "summary.json")) as f:
        nb = json.load(f)["notebook_printed"]
    assert abs(disc / (nb["EO_payoff_mean"] * np.exp(-0.08)) - 1) < 2e-6      # EO cell 8's printed payoff mean, discounted
    # what cell 20 prints as "Discounted Eq[V(T)]" is values[:, 0].mean() * exp(-r T), i.e. V0_RP discounted once more
    assert abs(log["V0_RP"] * np.exp(-0.08) - log["printed_discounted_V0"]) < 1e-3


def test_first_fit_follows_the_notebooks_log(eo_run):
    log, res, _, _ = eo_run
    h = res["history"][0]
    assert log["epochs_per_fit"][0] == 500                      # the notebook's first fit ran out of epochs
    assert h["epochs_run"] >= 350                               # patience 50 under the LR schedule: stops late or not at all
    assert abs(_last_loss(h) / log["fit_last_epoch_loss"][0] - 1) < 0.10


def test_later_fits_follow_the_notebooks_log(eo_run):
    log, res, _, _ = eo_run
    hist = res["history"]
    assert len(hist) == len(log["epochs_per_fit"]) == 52
    ours = np.array([_last_loss(h) for h in hist[1:]])
    theirs = np.array(log["fit_last_epoch_loss"][1:])
    # per-step losses fall from 2.9e-4 to 1.9e-4 along the backward sweep in both runs
    assert abs(ours[-10:].mean() / theirs[-10:].mean() - 1) < 0.10
    assert abs(ours[:10].mean() / theirs[:10].mean() - 1) < 0.10
    ep_ours, ep_theirs = sum(h["epochs_run"] for h in hist[1:]), sum(log["epochs_per_fit"][1:])
    assert abs(ep_ours / ep_theirs - 1) < 0.30, (ep_ours, ep_theirs)
    assert max(h["epochs_run"] for h in hist[1:]) < 100         # EarlyStopping(patience=7) fires in every later fit, as there


def test_price_and_hedge_at_t0_are_the_notebooks_within_the_seed_spread(eo_run):
    log, res, S0, disc = eo_run
    v0 = res["V0"] * S0
    assert abs(v0 / log["V0_RP"] - 1) < 0.15, (v0, log["V0_RP"])
    assert v0 > disc                                            # the method over-prices, there (8 %) and here
    assert abs(res["phi0"] - log["phi0"]) < 0.02 and abs(res["psi0"] - log["psi0"]) < 0.02
    assert abs(res["phi0"] + res["psi0"] - 1) < 1e-6            # psi = 1 - phi (EO cell 12)


def test_batch_order_identity_equals_index_order_and_a_permutation_changes_the_steps():
    """``batch_order`` (the step make-up handed to the oracle by the virtual-shard test of the CUDA path): the identity
    order is the default index order bit for bit; the virtual-shard order visits the same samples per epoch in other
    steps, so the trajectory differs while the first epoch's total sample set does not."""
    rng = np.random.default_rng(4)
    n, D = 4096, 2
    z = rng.standard_normal((n, D))
    S = np.empty((n, D + 1))
    S[:, 0] = 100.0
    for t in range(D):
        S[:, t + 1] = S[:, t] * np.exp((0.05 - 0.02) / 12 + 0.2 * np.sqrt(1 / 12) * z[:, t])
    bond = np.exp(0.05 * np.arange(D + 1) / 12)
    y = np.maximum(S[:, -1] - 100.0, 0.0) / 100.0
    kw = dict(net="eo", scale=100.0, epochs_first=3, epochs_next=2, patience_first=3, patience_next=2, batch_size=n // 8,
              shuffle_seed=None)
    a = ofit.backward_replication(S, bond, y, **kw)
    b = ofit.backward_replication(S, bond, y, batch_order=np.arange(n), **kw)
    assert a["V0"] == b["V0"] and np.array_equal(a["params"], b["params"])
    V, spe = 8, 8
    seg = n // V
    per = seg // spe
    order = np.concatenate([np.arange(v * seg + k * per, v * seg + (k + 1) * per) for k in range(spe) for v in range(V)])
    assert np.array_equal(np.sort(order), np.arange(n))
    c = ofit.backward_replication(S, bond, y, batch_order=order, **kw)
    assert not np.array_equal(a["params"], c["params"]) and abs(c["V0"] / a["V0"] - 1) < 0.05
