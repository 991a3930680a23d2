"""Host-side logic of the product layer that needs no GPU."""
import os

import numpy as np
import pytest

from oracle import fit as ofit, paths as opaths


def test_brownian_motion_dropin_matches_reference(golden_dir):
    import brownian_motion as bm
    g = np.load(os.path.join(golden_dir, "brownian.npz"))
    assert np.array_equal(bm.get_dW(5, 1), g["dW_5_1"]) and np.array_equal(bm.get_W(5, 1), g["W_5_1"])
    assert np.array_equal(bm.get_dW(64, 1234), g["dW_64_1234"]) and np.array_equal(bm.get_W(64, 1234), g["W_64_1234"])
    np.testing.assert_allclose(bm.get_W(5, 1), [0, 1.62434536, 1.01258895, 0.4844172, -0.58855142], atol=5e-9)
    assert bm.get_W(1, 3).tolist() == [0.0]


@pytest.mark.parametrize("T,dt,expect", [(1, 1 / 252, 253), (1, 1 / 52, 53), (1, 1 / 365, 366), (10, 1 / 365, 3651),
                                        (10, 1 / 100, 1001), (10, 1 / 12, 121)])
def test_grid_sizes(T, dt, expect):
    from orp_b200 import n_time_steps_of
    assert n_time_steps_of(T, dt) == expect == opaths.n_time_steps_of(T, dt)


def test_reduction_and_bond_match_oracle():
    from orp_b200 import bond_curve, reduction_of
    for n, T, reb in ((366, 1.0, 1 / 52), (1001, 10, 1 / 4), (121, 10, 10), (3651, 10, 1 / 4)):
        red, cols, n_red = opaths.reduction_of(n, T, reb)
        assert reduction_of(n, T, reb) == red
        b = bond_curve(0.03, T, n, red)
        assert b.shape == (n_red,) and np.array_equal(b, opaths.bond_curve(0.03, T, n)[cols])


def test_shard_ranges_tile_the_index_set():
    from orp_b200 import shard_range
    for total in (2 ** 16, 1000, 7):
        for world in (1, 2, 4, 8):
            got = [shard_range(total, r, world) for r in range(world)]
            assert got[0][0] == 0 and sum(c for _, c in got) == total
            for (s0, c0), (s1, _) in zip(got, got[1:]):
                assert s0 + c0 == s1


def test_init_params_layout_matches_oracle():
    from orp_b200 import init_params
    for net, bias in (("eo", 0.11), ("rp", [0.9, 0.1])):
        a, b = init_params(net, 1234, bias), ofit.init_params(net, 1234, bias)
        assert a.dtype == np.float32 and np.array_equal(a, b) and a.size == ofit.n_params(net)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import orp_b200
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        orp_b200.sobol_uint32(4, 3, 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        orp_b200.simulate_paths("gbm_log", dict(S0=1, mu=0, sigma=.1, dt=.1), m=4, steps=3)


def test_product_never_imports_the_oracle():
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                        "option-replicating-portfolio-with-neural-networks_b200")
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f


def test_cir_calibration_matches_the_reference_regression():
    """estimate_CIR_params against the reference's formulation (scikit-learn's
    LinearRegression(fit_intercept=False), Extra notebook cell 4) on a synthetic volatility series, and the
    CIRParams validity check (cell 3)."""
    import pytest
    from sklearn.linear_model import LinearRegression
    from orp_b200 import CIRParams, estimate_CIR_params
    rng = np.random.default_rng(3)
    a, b, c, n = 0.02, 0.16, 0.004, 4000
    s = np.empty(n)
    s[0] = 0.15
    for t in range(1, n):                                  # the reference's own recursion (RP.py:285, dt folded in)
        s[t] = s[t - 1] + a * (b - s[t - 1]) + c * np.sqrt(s[t - 1]) * rng.standard_normal()
    got = estimate_CIR_params(s)
    sq = np.sqrt(s[:-1])
    y, X = np.diff(s) / sq, np.stack([1.0 / sq, sq], axis=1)
    reg = LinearRegression(fit_intercept=False).fit(X, y)
    ref_a = -reg.coef_[1]
    ref = (ref_a, reg.coef_[0] / ref_a, np.std(y - reg.predict(X)))
    np.testing.assert_allclose([got.a, got.b, got.c], ref, rtol=1e-9)
    assert abs(got.b - b) < 0.02 and abs(got.c / c - 1) < 0.05
    with pytest.raises(ValueError, match="2ab"):
        CIRParams(0.001, 0.1, 0.1)
