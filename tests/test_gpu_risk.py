"""GPU tests of the per-step risk metrics: device quantiles by radix select against
``np.quantile`` — the call the reference makes on its per-path ``VaR`` array (``Replicating_Portfolio.py:122``).

Order statistics are selected exactly (integer keys), the interpolation is NumPy's ``_lerp`` in fp64: results must
agree to 1e-12 relative with ``np.quantile`` of the same fp32 values promoted to fp64."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

QS = [0.0, 0.02, 0.5, 0.98, 0.985, 0.99, 0.995, 1.0]


def _check(x, qs=QS):
    import orp_b200
    got = orp_b200.quantiles(torch.from_numpy(x).cuda(), qs)
    ref = np.quantile(x.astype(np.float64), qs)
    np.testing.assert_allclose(got, ref, rtol=1e-12, atol=0)


@pytest.mark.parametrize("n", [1, 2, 3, 257, 4099, 1_000_003])
def test_quantiles_match_numpy(cuda, n):
    rng = np.random.default_rng(n)
    _check((rng.standard_normal(n) * 3).astype(np.float32))


def test_quantiles_ties_signs_and_infinities(cuda):
    rng = np.random.default_rng(5)
    x = np.concatenate([np.zeros(500), -np.zeros(500), np.full(300, 1.5), rng.integers(-3, 4, 5000).astype(np.float64),
                        [1e-45, -1e-45, 3e38, -3e38]]).astype(np.float32)
    rng.shuffle(x)
    _check(x)
    # infinities: NumPy's own lerp turns (b - a) * g into NaN next to an infinity; the device path mirrors it
    y = np.concatenate([x, [np.inf, np.inf, -np.inf]]).astype(np.float32)
    qs = [0.0, 1e-5, 0.5, 0.99999, 1.0]
    got = __import__("orp_b200").quantiles(torch.from_numpy(y).cuda(), qs)
    with np.errstate(invalid="ignore"):
        ref = np.quantile(y.astype(np.float64), qs)
    np.testing.assert_array_equal(got, ref)


def test_quantiles_nan_and_errors(cuda):
    import orp_b200
    x = torch.tensor([1.0, float("nan"), 3.0], device=cuda)
    assert np.isnan(orp_b200.quantiles(x, [0.5, 0.99])).all()            # like np.quantile
    with pytest.raises(ValueError, match="range"):
        orp_b200.quantiles(torch.ones(4, device=cuda), [1.5])
    with pytest.raises(ValueError):
        orp_b200.quantiles(torch.ones(4, device=cuda, dtype=torch.float64), [0.5])


def test_quantiles_ignore_nan_is_nanquantile(cuda):
    """ignore_nan=True (used by fit_replication(skip_nonfinite=True)): NaN entries take no part, like np.nanquantile."""
    import orp_b200
    rng = np.random.default_rng(11)
    x = (rng.standard_normal(100_003) * 2).astype(np.float32)
    x[rng.choice(x.size, 37, replace=False)] = np.nan
    x[:3] = [np.nan, 5.0, np.nan]
    got = orp_b200.quantiles(torch.from_numpy(x).cuda(), QS, ignore_nan=True)
    np.testing.assert_allclose(got, np.nanquantile(x.astype(np.float64), QS), rtol=1e-12, atol=0)
    assert np.isnan(orp_b200.quantiles(torch.from_numpy(x).cuda(), QS)).all()
    allnan = torch.full((5,), float("nan"), device=cuda)
    assert np.isnan(orp_b200.quantiles(allnan, [0.5], ignore_nan=True)).all()


def test_step_var_quantiles_are_the_quantiles_of_the_residuals(cuda):
    """fit_replication(var_quantiles=...) reports, per time step, np.quantile of  values[:, t+1] - V(X1)  (the
    reference's VaR array) with the weights left by that step's fits."""
    import orp_b200
    from orp_b200.fit import ReplicationNet
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 4)
    tm = orp_b200.simulate_paths("gbm_log", p, m=13, steps=4, device=cuda, time_major=True)
    vals = torch.empty(2 ** 13, dtype=torch.float32, device=cuda)
    orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
    res = orp_b200.fit_replication(tm, orp_b200.bond_curve(0.05, 1.0, 5), vals, net="eo", scale=100.0,
                                   epochs_first=30, epochs_next=5, var_quantiles=(.98, .99))
    assert all(len(st["VaR"]) == 2 and st["VaR"][0] <= st["VaR"][1] for st in res["steps"])
    # last step (t_i = 0): recompute the residuals with the returned weights
    model = ReplicationNet("eo", res["params"], cuda)
    # values[:, 1] is not returned; check the terminal step instead by refitting nothing: residuals of a fresh model
    b = model.batch(tm[3], tm[4], vals, 1 / 100.0, float(orp_b200.bond_curve(0.05, 1.0, 5)[3]),
                    float(orp_b200.bond_curve(0.05, 1.0, 5)[4]))
    r = model.residuals(b, torch.empty_like(vals))
    got = orp_b200.quantiles(r, (.98, .99))
    np.testing.assert_allclose(got, np.quantile(r.cpu().numpy().astype(np.float64), (.98, .99)), rtol=1e-12)
    # and against the network evaluated by the oracle on the host
    from oracle import fit as ofit
    S = tm.t().cpu().numpy()
    bond = orp_b200.bond_curve(0.05, 1.0, 5)
    st = torch.from_numpy((S[:, 3] / 100).astype(np.float32))
    st1 = torch.from_numpy((S[:, 4] / 100).astype(np.float32))
    p1 = torch.stack([st1, torch.full_like(st, np.float32(bond[4] / 100))], dim=1)
    v1, _, _ = ofit.forward(torch.from_numpy(res["params"].copy()), "eo", st[:, None], p1)
    ref = vals.cpu().numpy() - v1.numpy()
    np.testing.assert_allclose(r.cpu().numpy(), ref, atol=2e-6)


def test_device_survivors_match_the_reference_draws_in_distribution(cuda):
    """SURVEY.md §8f rank 2.  The reference thins each path's population with NumPy's legacy binomial stream
    (Replicating_Portfolio.py:78-84) — sequential, not reproducible in parallel.  The device kernel draws the same
    binomial model from a counter-based generator: compare the moments of N_T over 2^14 paths of the notebook's
    mortality model (10 years, dt = 1/100) with the host draws.  Tolerances: mean within 6 (two independent sample
    means of std ~132 over 16384 paths differ by ~1.5), std within 5 %."""
    import orp_b200
    N, dt, steps, m = 10_000, 1 / 100, 1000, 14
    L = orp_b200.simulate_paths("linear", {"S0": 0.01, "mu": 0.075, "sigma": 0.000597, "dt": dt}, m=m, steps=steps,
                                seed=1234, device=cuda, time_major=True)
    frac = orp_b200.survivors(L, N, dt, keep_every=25)
    assert frac.shape == (41, 2 ** m) and bool((frac[0] == 1).all())
    f = frac.cpu().numpy().astype(np.float64)
    assert (np.diff(f, axis=0) <= 0).all() and (f >= 0).all()
    Lh = L.cpu().numpy()
    Nh = np.full(2 ** m, N)
    for t in range(1, steps + 1):
        np.random.seed(1234 + t)
        Nh = np.random.binomial(Nh, np.exp(-Lh[t] * dt))
    n_T = f[-1] * N
    assert abs(n_T.mean() - Nh.mean()) < 6.0, (n_T.mean(), Nh.mean())
    assert abs(n_T.std() / Nh.std() - 1) < 0.05, (n_T.std(), Nh.std())
    # the survivors are integers, and a shard draws exactly what the whole set draws for the same paths
    assert np.abs(n_T - np.rint(n_T)).max() < 1e-2
    part = orp_b200.survivors(L[:, 5000:9000].contiguous(), N, dt, keep_every=25, start=5000)
    assert torch.equal(part, frac[:, 5000:9000])


def test_device_survivors_one_step_binomial_moments(cuda):
    """One step with a constant intensity: deaths ~ Binomial(N, q); mean N q and variance N q (1 - q) over 2^18 paths,
    for a rare-event q (the reference's regime: inversion from 0), for large means (n q = 130, 1000 and 7e5: the exact
    transformed-rejection sampler — the inversion's P(D = 0) underflows there and np.random.binomial handles any n) and
    for q > 1/2 (drawn through the complement, both samplers)."""
    import orp_b200
    n_paths = 2 ** 18
    for N, lam in ((10_000, 0.015), (500, 30.0), (1_000_000, 0.1), (1_000_000, 120.0), (2000, 300.0), (50, 200.0)):
        dt = 0.01
        L = torch.full((2, n_paths), lam, dtype=torch.float64, device=cuda)
        frac = orp_b200.survivors(L, N, dt, seed=7)
        deaths = (1.0 - frac[1].double()) * N
        q = 1 - np.exp(-lam * dt)
        mean, var = float(deaths.mean()), float(deaths.var())
        assert abs(mean - N * q) < 5 * np.sqrt(N * q * (1 - q) / n_paths) + 1e-3 * N * q, (mean, N * q)
        assert abs(var / (N * q * (1 - q)) - 1) < 0.03, (var, N * q * (1 - q))


def test_host_picked_and_device_picked_selection_agree(cuda):
    """The two forms of the C ABI: orp_select_hist / orp_select_next with the digit picked by the caller between
    passes, and orp_select_hist_multi / orp_select_pick / orp_select_next_multi with the state on the device."""
    from orp_b200 import _lib
    from orp_b200.risk import _select, BINS
    L = _lib.lib()
    rng = np.random.default_rng(3)
    x = torch.from_numpy(np.round(rng.standard_normal(200_001) * 50).astype(np.float32) / 8 + 0.0625).to(cuda)     # many ties, no +-0
    ks = np.array([0, 1234, 100_000, 200_000])
    keys, le, nxt, nan, bad = _select(x, ks, None)
    assert not nan.any() and not bad.any()
    srt = np.sort(x.cpu().numpy())
    for j, k in enumerate(ks):
        hist = torch.zeros(BINS, dtype=torch.int64, device=cuda)
        prefix = mask = 0
        kk = int(k)
        for shift in (24, 16, 8, 0):
            hist.zero_()
            _lib.check(L.orp_select_hist(_lib.ptr(x), x.numel(), shift, prefix, mask, _lib.ptr(hist), _lib.stream_ptr()))
            cum = np.cumsum(hist.cpu().numpy()[:256])
            d = int(np.searchsorted(cum, kk, side="right"))
            kk -= int(cum[d - 1]) if d else 0
            prefix |= d << shift
            mask |= 0xFF << shift
        out = torch.tensor([0, 0xFFFFFFFF], dtype=torch.int64, device=cuda)
        _lib.check(L.orp_select_next(_lib.ptr(x), x.numel(), prefix, _lib.ptr(out), _lib.stream_ptr()))
        o = out.cpu().numpy()
        assert (prefix, int(o[0]), int(o[1])) == (int(keys[j]), int(le[j]), int(nxt[j]))
        from orp_b200.risk import _key_to_float
        assert _key_to_float(prefix) == srt[k] and int(o[0]) == int((srt <= srt[k]).sum())
    # k beyond the element count is reported, not silently clamped
    assert _select(x, np.array([x.numel()]), None)[4].all()
