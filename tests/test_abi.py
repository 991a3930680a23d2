"""CPU-side checks of the C-ABI boundary: the library loads, exports every symbol the header declares
(and nothing the binding does not know), and its host-only integer routines agree with the oracle.
No compute call touches a GPU here."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import bridge, sobol as osobol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "orp_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(orp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from orp_b200 import _lib
    handle = ctypes.CDLL(_lib.LIB_PATH) if os.path.exists(_lib.LIB_PATH) else _lib.lib()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(handle, n), f"{n} declared in include/orp_b200.h but not exported"
    assert sorted(_lib.PROTOTYPES) == names, "ctypes prototypes out of step with the header"


def test_struct_layouts_match_header():
    from orp_b200 import _lib
    assert ctypes.sizeof(_lib.PathParams) == 6 * 4 + 8 * 8
    assert ctypes.sizeof(_lib.FitBatch) == 4 * 4 + 8 + 5 * 8 + 3 * 8 + 2 * 4 + 8
    assert _lib.lib().orp_fit_n_params(0) == 97 and _lib.lib().orp_fit_n_params(1) == 122   # Keras summaries
    assert _lib.lib().orp_version() >= 100


@pytest.mark.parametrize("d,seed", [(1, 7), (2, 1235), (53, 1235), (253, 1234), (253, 1235), (366, 1235),
                                    (3651, 1235), (21201, 3)])          # RP.py at T = 10, dt = 1/365; SciPy's maximum
def test_host_setup_bit_exact(d, seed):
    from orp_b200.sobol import SobolEngine
    e = SobolEngine(d, seed)
    ref = osobol.scipy_engine(d, seed)
    assert np.array_equal(e.sv, ref._sv) and np.array_equal(e.shift, ref._shift)
    assert np.array_equal(e.sv32[:30, :d].T, ref._sv) and np.array_equal(e.sv32[31, :d], ref._shift)
    assert not e.sv32[30].any() and not e.sv32[:, d:].any()
    # lane table against the closed form: index i = 32 p + l of warp tile 0 (shift and base removed)
    x = osobol.points_u32(ref._sv, np.zeros(d, dtype=np.uint32), 0, 64)          # (64, d)
    assert np.array_equal(e.tlane[:, :, 0], x[:32].T) and np.array_equal(e.tlane[:, :, 1], x[32:].T)


def test_unscrambled_direction_numbers():
    from orp_b200.sobol import SobolEngine
    from scipy.stats import qmc
    e = SobolEngine(300, None, scramble=False)
    assert np.array_equal(e.sv, qmc.Sobol(300, scramble=False)._sv)


def test_errors_follow_the_reference_convention():
    from orp_b200 import _lib
    from orp_b200.sobol import SobolEngine
    with pytest.raises(ValueError, match="Maximum supported dimensionality"):
        SobolEngine(21202, 1)
    L = _lib.lib()
    v = np.zeros((1, 30), dtype=np.uint32)
    rc = L.orp_sobol_direction_numbers(None, None, 1, 30, _lib.ptr(v))
    assert rc == _lib.ORP_E_VALUE and b"null" in L.orp_last_error()
    with pytest.raises(ValueError):
        _lib.check(rc)
    assert L.orp_bridge_schedule(0, _lib.ptr(v)) == _lib.ORP_E_VALUE


@pytest.mark.parametrize("D", [1, 2, 5, 52, 252, 1000])
def test_bridge_program_matches_oracle_schedule(D):
    from orp_b200 import bridge_schedule_host
    ops = bridge_schedule_host(D)
    node, left, right, dim = bridge.bridge_schedule(D)
    wl, wr, sd = bridge.bridge_weights(D)
    order = np.argsort(ops["dim"])
    assert np.array_equal(ops["dim"][order], dim)
    np.testing.assert_allclose(ops["wl"][order], wl, rtol=0, atol=1e-15)
    np.testing.assert_allclose(ops["wr"][order], wr, rtol=0, atol=1e-15)
    np.testing.assert_allclose(ops["sd"][order], sd, rtol=1e-15)
    assert ops["emits"].sum() == D
    # replay the depth-first program on the host: it must emit W_1..W_D of the breadth-first oracle
    rng = np.random.default_rng(D)
    Z = rng.standard_normal((3, D + 1))
    ref = np.cumsum(bridge.bridge_normals(Z)[:, 1:], axis=1)
    stack, cur, emitted = [], np.zeros(3), []
    for op in ops:
        right_v = stack[-1] if stack else np.zeros(3)
        stack.append(op["wl"] * cur + op["wr"] * right_v + op["sd"] * Z[:, op["dim"]])
        for _ in range(op["emits"]):
            cur = stack.pop()
            emitted.append(cur)
    np.testing.assert_allclose(np.stack(emitted, axis=1), ref, rtol=1e-12, atol=1e-12)
