"""The oracle's Sobol restatement against the installed SciPy and the committed golden points."""
import os

import numpy as np
import pytest

from oracle import sobol


@pytest.mark.parametrize("d,seed", [(1, 1234), (2, 1235), (5, 1235), (53, 1235), (253, 1234), (253, 1235)])
def test_setup_equals_scipy_private_state(d, seed):
    sv, shift = sobol.sobol_setup(d, seed)
    eng = sobol.scipy_engine(d, seed)
    assert sv.dtype == np.uint32 and np.array_equal(sv, eng._sv)
    assert np.array_equal(shift, eng._shift)


@pytest.mark.parametrize("d,seed,start,count", [(7, 1234, 0, 256), (53, 1235, 0, 1024), (53, 1235, 1000, 77),
                                                (253, 1235, 2 ** 22 - 64, 64), (253, 1235, 123456, 100)])
def test_closed_form_points_equal_scipy(d, seed, start, count):
    sv, shift = sobol.sobol_setup(d, seed)
    assert np.array_equal(sobol.points_u32(sv, shift, start, count), sobol.scipy_points_u32(d, seed, start, count))


def test_random_base2_equals_window(golden_dir):
    u = sobol.scipy_engine(9, 1235).random_base2(8)
    sv, shift = sobol.sobol_setup(9, 1235)
    assert np.array_equal((u * 2.0 ** 30).astype(np.uint32), sobol.points_u32(sv, shift, 0, 256))


def test_golden_points(golden_dir):
    g = np.load(os.path.join(golden_dir, "sobol_points.npz"))
    for key in g.files:
        if not key.startswith("pts_"):
            continue
        _, d, seed, start = key.split("_")
        d, seed, start = int(d), int(seed), int(start)
        cols = g[f"cols_{d}_{seed}_{start}"]
        sv, shift = g[f"sv_{d}_{seed}"], g[f"shift_{d}_{seed}"]
        assert np.array_equal(sobol.points_u32(sv, shift, start, 64)[:, cols], g[key]), key


def test_zero_coordinate_gives_minus_inf():
    # Sobol(253, seed=1235) path 2,631,230, dimension 236 is exactly 0.0
    w = sobol.normals_window(253, 1235, 2_631_230, 1)
    assert w[0, 236] == -np.inf
    sv, shift = sobol.sobol_setup(253, 1235)
    assert sobol.points_u32(sv, shift, 2_631_230, 1)[0, 236] == 0
