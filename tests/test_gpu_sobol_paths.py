"""GPU parity tests (through the C ABI): Sobol points bit-exact, inverse normal, paths, payoff sums.
Tolerances: integer points bit-exact; normals 2e-11 relative-or-absolute against scipy.special.ndtri (the 32-byte-row
table is good to 9e-12, five orders inside what the paths need); fp64 path values 1e-6 relative (north_star;
observed ~1e-11), fp32 1e-4; sums 1e-9 relative."""
import json
import os

import numpy as np
import pytest

from oracle import paths as opaths, sobol as osobol

pytestmark = pytest.mark.gpu

RTOL64, RTOL32 = 1e-6, 1e-4


def _rel_inf_aware(got, ref, rtol):
    """Relative agreement with identical placement of +-inf / NaN."""
    fin = np.isfinite(ref)
    assert np.array_equal(fin, np.isfinite(got)), "non-finite entries differ"
    assert np.array_equal(np.isnan(ref), np.isnan(got))
    if (~fin & ~np.isnan(ref)).any():
        assert np.array_equal(ref[~fin & ~np.isnan(ref)], got[~fin & ~np.isnan(ref)])
    np.testing.assert_allclose(got[fin], ref[fin], rtol=rtol, atol=0)


@pytest.mark.parametrize("d,seed,start,count", [
    (2, 1235, 0, 2 ** 16),                 # C1 shape
    (53, 1235, 0, 4096), (53, 1235, 2 ** 20 - 100, 100),          # C2
    (253, 1235, 0, 1024), (253, 1234, 2 ** 22 - 64, 64),          # C3
    (253, 1235, 2 ** 24 // 3, 257),                               # C4, ragged window
    (253, 1235, 2 ** 26 - 192, 192), (253, 1235, 2_631_230 - 5, 11),   # C5 end; the exact-zero coordinate
    (1, 7, 0, 100), (366, 1235, 12345, 300), (9, 1, 63, 2), (9, 1, 64, 1),
])
def test_points_bit_exact(cuda, d, seed, start, count):
    import orp_b200
    ref = osobol.scipy_points_u32(d, seed, start, count)
    for layout in ("points", "dims"):
        got = orp_b200.sobol_uint32(None, d, seed, start=start, count=count, layout=layout).cpu().numpy()
        got = got if layout == "points" else got.T
        assert got.dtype == np.uint32 and np.array_equal(got, ref), (layout, d, seed, start)


def test_points_random_base2_and_golden(cuda, golden_dir):
    import orp_b200
    u = osobol.scipy_engine(5, 1234).random_base2(10)
    x = orp_b200.sobol_uint32(10, 5, 1234).cpu().numpy()
    assert np.array_equal(x * 2.0 ** -30, u)
    g = np.load(os.path.join(golden_dir, "sobol_points.npz"))
    for key in (k for k in g.files if k.startswith("pts_")):
        _, d, seed, start = key.split("_")
        cols = g[f"cols_{d}_{seed}_{start}"]
        got = orp_b200.sobol_uint32(None, int(d), int(seed), start=int(start), count=64).cpu().numpy()
        assert np.array_equal(got[:, cols], g[key]), key


def test_points_empty_and_errors(cuda):
    import orp_b200
    assert orp_b200.sobol_uint32(None, 3, 1, start=5, count=0).shape == (0, 3)
    with pytest.raises(ValueError, match="distinct points"):
        orp_b200.sobol_uint32(31, 3, 1)
    with pytest.raises(ValueError, match="dimensionality"):
        orp_b200.sobol_uint32(4, 21202, 1)


def test_ndtri_matches_scipy(cuda):
    import torch
    import scipy.special as sp
    import orp_b200
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.integers(0, 2 ** 30, 1_000_000), np.arange(0, 4096), 2 ** 30 - 1 - np.arange(0, 4096),
                        2 ** 29 + np.arange(-2048, 2048), (np.arange(1, 8)[:, None] * 2 ** 27 + np.arange(-3, 4)).ravel(),
                        2 ** np.arange(0, 30), 2 ** 30 - 2 ** np.arange(0, 30)]).astype(np.uint32)
    z = orp_b200.ndtri_u32(torch.from_numpy(x.view(np.int32)).to(cuda)).cpu().numpy()
    ref = sp.ndtri(x.astype(np.float64) * 2.0 ** -30)
    assert z[x == 0][0] == -np.inf and np.all(z[x == 2 ** 29] == 0.0)
    fin = np.isfinite(ref)
    assert np.array_equal(fin, np.isfinite(z))
    err = np.abs(z[fin] - ref[fin]) / np.maximum(1.0, np.abs(ref[fin]))
    assert err.max() < 2e-11, err.max()      # table: 9e-12 (degree 5, c2..c5 in fp32: tools/gen_ndtri_table.py) + scipy: 4e-16


def test_sobol_normal_matches_reference_call_form(cuda):
    import orp_b200
    w = orp_b200.sobol_normal(8, 13, 1235).cpu().numpy()
    ref = osobol.sobol_norm(8, 13, 1235)
    np.testing.assert_allclose(w, ref, rtol=2e-11, atol=2e-11)


SIMS = {
    "gbm_log": dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252),
    "gbm_arith": dict(S0=1.0, mu=0.09464, sigma=0.15965, dt=1 / 100),
    "sv": dict(S0=1.0, mu=0.03, s0=0.16679, a=0.0033281299103885727, b=0.1562947229160206, c=0.01583113545983743, dt=1 / 252),
    "linear": dict(S0=0.01, mu=0.075, sigma=0.000597, dt=1 / 100),
}


def _oracle_paths(model, p, d, start, count, ordering="incremental"):
    if model == "linear":
        W = osobol.normals_window(d, 1234, start, count)
        L = np.empty((count, d))
        L[:, 0] = p["S0"]
        for t in range(1, d):        # Replicating_Portfolio.py:75-76
            L[:, t] = L[:, t - 1] + (p["mu"] * L[:, t - 1] * p["dt"] + p["sigma"] * np.sqrt(p["dt"]) * W[:, t])
        return L
    return opaths.simulate_window(model, p, d, start, count, ordering=ordering)


@pytest.mark.parametrize("model", ["gbm_log", "gbm_arith", "sv", "linear"])
@pytest.mark.parametrize("steps,start,count,keep", [(252, 0, 2048, 1), (52, 1000, 333, 1), (252, 2 ** 22 - 128, 128, 21),
                                                   (1, 0, 4096, 1), (100, 77, 1, 7), (365, 4096, 640, 7)])
def test_paths_match_oracle(cuda, model, steps, start, count, keep):
    import orp_b200
    p = SIMS[model]
    seed = 1234 if model == "linear" else 1235
    got = orp_b200.simulate_paths(model, p, steps=steps, start=start, count=count, keep_every=keep, seed=seed)
    assert tuple(got.shape) == (count, steps // keep + 1) and got.stride(0) == 1
    ref = _oracle_paths(model, p, steps + 1, start, count)[:, ::keep]
    _rel_inf_aware(got.cpu().numpy(), ref, RTOL64)
    got32 = orp_b200.simulate_paths(model, p, steps=steps, start=start, count=count, keep_every=keep, seed=seed,
                                    dtype="float32").cpu().numpy()
    _rel_inf_aware(got32.astype(np.float64), ref.astype(np.float32).astype(np.float64), RTOL32)


@pytest.mark.parametrize("model", ["gbm_log", "gbm_arith", "sv"])
def test_exact_zero_coordinate_propagates(cuda, model):
    """Path 2,631,230 of Sobol(253, seed 1235) holds u = 0 in dimension 236:
    log-Euler -> exp(-inf) = 0 from there on, arithmetic Euler -> -inf then NaN."""
    import orp_b200
    p = dict(SIMS[model], dt=1 / 252)
    got = orp_b200.simulate_paths(model, p, steps=252, start=2_631_230 - 3, count=8).cpu().numpy()
    ref = _oracle_paths(model, p, 253, 2_631_230 - 3, 8)
    assert not np.isfinite(ref[3]).all() or (ref[3, 236:] == 0).all()
    _rel_inf_aware(got, ref, RTOL64)


@pytest.mark.parametrize("model", ["gbm_log", "gbm_arith", "sv"])
@pytest.mark.parametrize("steps,count", [(52, 1024), (252, 512), (7, 100), (1, 64)])
def test_bridge_paths_match_oracle(cuda, model, steps, count):
    import orp_b200
    p = dict(SIMS[model], dt=1.0 / steps)
    got = orp_b200.simulate_paths(model, p, steps=steps, start=0, count=count, ordering="bridge").cpu().numpy()
    ref = _oracle_paths(model, p, steps + 1, 0, count, ordering="bridge")
    _rel_inf_aware(got, ref, RTOL64)


def test_paths_match_reference_fixtures(cuda, golden_dir):
    """Against arrays produced by the unmodified reference (tests/golden/make_golden.py)."""
    import orp_b200
    g = np.load(os.path.join(golden_dir, "rp_paths.npz"))
    p = json.loads(str(g["params"]))
    Y = orp_b200.simulate_paths("gbm_arith", dict(S0=p["Y"], mu=p["mu"], sigma=p["sigma"], dt=p["dt"]),
                                m=p["n_paths"], steps=1000, keep_every=int(g["reduction"])).cpu().numpy()
    np.testing.assert_allclose(Y, g["Y_reduced"], rtol=RTOL64)
    L = orp_b200.simulate_paths("linear", dict(S0=p["l0"], mu=p["c"], sigma=p["ita"], dt=p["dt"]), m=p["n_paths"],
                                steps=1000, keep_every=125, seed=1234).cpu().numpy()
    np.testing.assert_allclose(L, g["L_cols"], rtol=RTOL64)
    e = np.load(os.path.join(golden_dir, "eo_paths.npz"))
    S = orp_b200.simulate_paths("gbm_log", dict(S0=float(e["S0"]), mu=float(e["r"]), sigma=float(e["sigma"]), dt=1 / 365),
                                m=int(e["n_paths_log2"]), steps=365, keep_every=int(e["reduction"])).cpu().numpy()
    np.testing.assert_allclose(S, e["S_weekly"], rtol=RTOL64)
    s = np.load(os.path.join(golden_dir, "rp_sv_paths.npz"))
    q = json.loads(str(s["params"]))
    for tag, c in (("c075", 0.075), ("c0158", 0.01583113545983743)):
        sim = dict(S0=q["Y"], mu=q["mu"], s0=q["s0"], a=q["a"], b=q["b"], c=c, dt=q["dt"])
        Ysv = orp_b200.simulate_paths("sv", sim, m=q["n_paths"], steps=1000, keep_every=125).cpu().numpy()
        np.testing.assert_allclose(Ysv, s[f"{tag}_Y_full_cols"], rtol=RTOL64)


@pytest.mark.parametrize("kind", ["call", "put", "floor"])
def test_payoff_sums_match_oracle(cuda, kind):
    import orp_b200
    p = SIMS["gbm_log"]
    strikes = np.repeat(np.linspace(80, 120, 8), 8)
    cols = np.tile(np.arange(31, 253, 31)[:8], 8)               # C4: 8 strikes x 8 maturities
    tm = orp_b200.simulate_paths("gbm_log", p, steps=252, start=512, count=3000, time_major=True)
    sums, otm = orp_b200.payoff_sums(tm, strikes, cols, kind, s0=p["S0"])
    rs, ro = opaths.payoff_sums_window("gbm_log", p, 253, 512, 3000, strikes, cols, kind)
    np.testing.assert_allclose(sums.cpu().numpy(), rs, rtol=1e-9)
    assert np.array_equal(otm.cpu().numpy(), ro)
    # terminal values written for the backward loop (fp32 cast of payoff / S0)
    import torch
    vals = torch.empty(3000, dtype=torch.float32, device=cuda)
    orp_b200.payoff_sums(tm, [100.0], [-1], kind, s0=p["S0"], values_out=vals, value_scale=1 / p["S0"])
    ref = opaths.payoff(tm[-1].cpu().numpy(), 100.0, kind) / p["S0"]
    np.testing.assert_allclose(vals.cpu().numpy(), ref.astype(np.float32), rtol=1e-6)


def test_payoff_nan_and_inf_semantics(cuda):
    """np.maximum propagates NaN, np.where(NaN > K, ., K) gives K; counts follow EO cell 8 / RP.py:89."""
    import torch
    import orp_b200
    col = np.array([np.nan, -np.inf, np.inf, 0.0, 150.0, 100.0])
    tm = torch.from_numpy(col[None, :].copy()).to(cuda)
    for kind in ("call", "put", "floor"):
        sums, otm = orp_b200.payoff_sums(tm, [100.0], [0], kind, s0=100.0)
        pay = opaths.payoff(col, 100.0, kind)
        ref_otm = np.count_nonzero(col < 100.0) if kind == "floor" else np.count_nonzero(~(pay > 0))
        got = float(sums[0])
        assert (np.isnan(got) and np.isnan(pay.sum())) or got == pay.sum(), kind
        assert float(otm[0]) == ref_otm, kind


def test_large_run_properties(cuda):
    """Size-independent checks at a size the oracle cannot materialise quickly (2^22 x 52, C2-like):
    shards tile the set (linearity), relaunch is bit-identical (idempotence), put-call parity, martingale."""
    import orp_b200
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 52)
    m, D = 22, 52
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True)
    tm2 = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, time_major=True)
    assert bool((tm == tm2).all())
    c, _ = orp_b200.payoff_sums(tm, [100.0], [-1], "call")
    q, _ = orp_b200.payoff_sums(tm, [100.0], [-1], "put")
    n = 2 ** m
    mean_T = float(tm[-1].sum()) / n
    assert abs((float(c[0]) - float(q[0])) / n - (mean_T - 100.0)) < 1e-9
    assert abs(mean_T / (100.0 * np.exp(0.05)) - 1) < 1e-5
    price = np.exp(-0.05) * float(c[0]) / n
    assert abs(price - opaths.black_scholes_call(100, 100, 1, .05, .2)) < 0.11      # weekly Euler bias + QMC error
    parts = 0.0
    for r in range(4):
        sh = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, shard=(r, 4), time_major=True)
        assert bool((sh == tm[:, r * n // 4:(r + 1) * n // 4]).all())
        parts += float(orp_b200.payoff_sums(sh, [100.0], [-1], "call")[0][0])
    assert abs(parts / float(c[0]) - 1) < 1e-12
    # an oracle window in the middle of the big run
    ref = opaths.simulate_window("gbm_log", p, D + 1, 3_000_000, 256)
    np.testing.assert_allclose(tm[:, 3_000_000:3_000_256].t().cpu().numpy(), ref, rtol=RTOL64)


def test_c4_strike_maturity_sweep_64_options(cuda):
    """BASELINE config C4 in small: 64 European calls (8 strikes x 8 maturity columns) priced off ONE 252-step GBM
    Sobol set in one payoff call; sums against the oracle (1e-9), prices against Black-Scholes (Euler + QMC error)."""
    import orp_b200
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252)
    m, D = 14, 252
    strikes = np.repeat(np.linspace(80.0, 120.0, 8), 8)
    cols = np.tile(np.arange(1, 9) * 31 + 4, 8)                     # 35, 66, ..., 252
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, device=cuda, time_major=True)
    sums, otm = orp_b200.payoff_sums(tm, strikes, cols, "call")
    rs, ro = opaths.payoff_sums_window("gbm_log", p, D + 1, 0, 2 ** m, strikes, cols, "call")
    np.testing.assert_allclose(sums.cpu().numpy(), rs, rtol=1e-9)
    assert np.array_equal(otm.cpu().numpy(), ro)
    prices, p_otm = orp_b200.price_european(tm, strikes, cols, "call", r=0.05, dt=1 / 252)
    bs = np.array([opaths.black_scholes_call(100.0, k, c / 252, 0.05, 0.2) for k, c in zip(strikes, cols)])
    assert np.abs(prices - bs).max() < 0.1              # QMC error of 2^14 paths in the far out-of-the-money corner
    assert ((p_otm >= 0) & (p_otm <= 1)).all()


# ---- fused generate + payoff: the paths are never stored -------------------------------------------------------------

@pytest.mark.parametrize("model,params,D", [
    ("gbm_log", dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252), 252),              # table-fused log-Euler step (C4 / C5 model)
    ("gbm_log", dict(S0=100.0, mu=0.05, sigma=0.9, dt=0.25), 8),                    # large steps: the unfused log model
    ("gbm_euler", dict(S0=1.0, mu=0.08, sigma=0.15, dt=1 / 100), 120),              # RP.py:65
    ("sv", dict(S0=1.0, mu=0.03, s0=0.16679, a=0.0033281299103885727, b=0.1562947229160206, c=0.01583113545983743, dt=1 / 252), 60),
])
@pytest.mark.parametrize("kind", ["call", "put", "floor"])
def test_fused_generate_and_payoff_equals_stored_paths(cuda, model, params, D, kind):
    """orp_paths_payoff against orp_paths + orp_payoff_reduce on the same index window (start not on a tile boundary,
    ragged count): sums to 1e-12 (another summation order), out-of-the-money counts exactly; and against the oracle's
    NumPy payoff on the reference's path code to 1e-9."""
    import orp_b200
    from oracle import paths as opaths
    start, count = 37, 70_001
    s0 = params["S0"]
    ks = np.array([0.8, 0.9, 1.0, 1.0, 1.1, 1.2, 0.95, 1.05, 1.0, 1.3, 0.7]) * s0
    ts = np.array([D, D // 2, D, 1, D // 3, D, 0, D - 1, D // 2, D, D // 2])        # 3 share D // 2, 4 share D; t = 0 and t = 1
    tm = orp_b200.simulate_paths(model, params, steps=D, start=start, count=count, device=cuda, time_major=True)
    ref_s, ref_o = orp_b200.payoff_sums(tm, ks, ts, kind, s0=s0)
    got_s, got_o = orp_b200.payoff_sums_fused(model, params, ks, ts, kind, steps=D, start=start, count=count, device=cuda, s0=s0)
    np.testing.assert_allclose(got_s.cpu().numpy(), ref_s.cpu().numpy(), rtol=1e-12)
    assert np.array_equal(got_o.cpu().numpy(), ref_o.cpu().numpy())
    p_or = dict(params)
    osum, ootm = opaths.payoff_sums_window({"gbm_euler": "gbm_arith"}.get(model, model), p_or, D + 1, start, count, ks, ts, kind)
    np.testing.assert_allclose(got_s.cpu().numpy(), osum, rtol=1e-9)
    assert np.array_equal(got_o.cpu().numpy(), ootm)


def test_fused_sweep_c4_shape_and_errors(cuda):
    """BASELINE C4's shape on a smaller set: 64 calls = 8 strikes x 8 maturities off one 2^18-path 252-step GBM set, from the
    fused kernel, against the stored-path sweep (1e-12) and Black-Scholes; more than 8 strikes per maturity (several
    groups on one step); bad arguments are ValueError."""
    import orp_b200
    from oracle import paths as opaths
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252)
    m, D = 18, 252
    ks = np.repeat(np.linspace(80.0, 120.0, 8), 8)
    ts = np.tile(np.arange(1, 9) * (D // 8) + D % 8, 8)
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, device=cuda, time_major=True)
    ref, _ = orp_b200.price_european(tm, ks, ts, "call", r=0.05, dt=1 / 252)
    got, potm = orp_b200.price_european_fused("gbm_log", p, ks, ts, "call", r=0.05, m=m, steps=D, device=cuda)
    np.testing.assert_allclose(got, ref, rtol=1e-12)
    bs = np.array([opaths.black_scholes_call(100.0, k, t / 252, 0.05, 0.2) for k, t in zip(ks, ts)])
    assert np.abs(got - bs).max() < 2e-2 and (potm > 0).all() and (potm < 1).all()
    ks2, ts2 = np.linspace(60.0, 140.0, 20), np.full(20, D)                          # 20 strikes on one maturity: 3 groups
    ref2, _ = orp_b200.price_european(tm, ks2, ts2, "put", r=0.05, dt=1 / 252)
    got2, _ = orp_b200.price_european_fused("gbm_log", p, ks2, ts2, "put", r=0.05, m=m, steps=D, device=cuda)
    np.testing.assert_allclose(got2, ref2, rtol=1e-12)
    with pytest.raises(ValueError):
        orp_b200.payoff_sums_fused("gbm_log", p, [100.0], [D + 1], m=m, steps=D, device=cuda)
    with pytest.raises(ValueError):
        orp_b200.payoff_sums_fused("gbm_log", p, np.full(65, 100.0), np.full(65, D), m=m, steps=D, device=cuda)


@pytest.mark.parametrize("kind", ["call", "put", "floor"])
@pytest.mark.parametrize("dtype", ["float64", "float32"])
def test_strike_ladder_bucket_form_equals_the_pair_form_and_numpy(cuda, kind, dtype, monkeypatch):
    """Several options on ONE column go through payoff_bucket_kernel (strikes sorted per chunk, one accumulation per
    path into the bucket 'number of options in the money'): unsorted strikes with ties, a short last chunk, strikes
    outside the price range, NaN and +-inf prices — against NumPy (sums 1e-12, out-of-the-money counts exact) and against
    the pair form (ORP_B200_NO_BUCKETS)."""
    import torch
    import orp_b200
    rng = np.random.default_rng(17)
    n = 50_003
    col = (100.0 * np.exp(0.3 * rng.standard_normal(n))).astype(dtype)
    tm = torch.from_numpy(np.stack([np.full(n, 100.0, dtype=dtype), col])).to(cuda)
    strikes = np.array([120.0, 80.0, 100.0, 100.0, 95.5, 300.0, 1.0, 110.0, 90.0, 1e6, 100.25])      # 8 + 3
    cols = np.ones(strikes.size, dtype=np.int64)

    def numpy_ref(c):
        c = c.astype(np.float64)
        out_s, out_o = [], []
        for K in strikes:
            pay = opaths.payoff(c, K, kind)
            out_s.append(pay.sum())
            out_o.append(np.count_nonzero(c < 100.0) if kind == "floor" else np.count_nonzero(~(pay > 0)))
        return np.array(out_s), np.array(out_o, dtype=np.float64)

    sums, otm = orp_b200.payoff_sums(tm, strikes, cols, kind, s0=100.0)
    rs, ro = numpy_ref(col)
    np.testing.assert_allclose(sums.cpu().numpy(), rs, rtol=1e-12)
    assert np.array_equal(otm.cpu().numpy(), ro)
    vals = torch.empty(n, dtype=torch.float32, device=cuda)
    orp_b200.payoff_sums(tm, strikes, cols, kind, s0=100.0, values_out=vals, value_scale=0.01)
    np.testing.assert_allclose(vals.cpu().numpy(), (opaths.payoff(col.astype(np.float64), strikes[0], kind) * 0.01).astype(np.float32), rtol=1e-6)
    monkeypatch.setenv("ORP_B200_NO_BUCKETS", "1")
    sums2, otm2 = orp_b200.payoff_sums(tm, strikes, cols, kind, s0=100.0)
    monkeypatch.delenv("ORP_B200_NO_BUCKETS")
    np.testing.assert_allclose(sums.cpu().numpy(), sums2.cpu().numpy(), rtol=1e-12)
    assert np.array_equal(otm.cpu().numpy(), otm2.cpu().numpy())
    # non-finite prices: same rules as the single-option pass (np.maximum propagates NaN; np.where(NaN > K, ., K) = K)
    bad = col.copy()
    bad[[5, 77, 4000]] = [np.nan, np.inf, -np.inf]
    tm_bad = torch.from_numpy(np.stack([np.full(n, 100.0, dtype=dtype), bad])).to(cuda)
    sums3, otm3 = orp_b200.payoff_sums(tm_bad, strikes, cols, kind, s0=100.0)
    with np.errstate(invalid="ignore"):
        rs3, ro3 = numpy_ref(bad)
    got = sums3.cpu().numpy()
    assert np.array_equal(np.isnan(got), np.isnan(rs3)) and np.array_equal(np.isinf(got), np.isinf(rs3))
    fin = np.isfinite(rs3)
    np.testing.assert_allclose(got[fin], rs3[fin], rtol=1e-12)
    assert np.array_equal(otm3.cpu().numpy(), ro3)
