"""Oracle: scrambled Sobol points exactly as ``scipy.stats.qmc.Sobol(d, scramble=True,
seed=seed)`` produces them (TEST INFRASTRUCTURE — see ``oracle/__init__.py``).

The reference obtains its normals through ``sobol_norm`` (``Replicating_Portfolio.py:54-57``,
``European Options.ipynb`` cell 4) which is SciPy's engine with the default ``bits=30``.
SciPy's ``_sobol`` extension ships only as a compiled module, so the algorithm is
restated here from its documented behaviour (Joe & Kuo direction numbers, Matousek
LMS + digital shift, Gray-code generation; ``scipy/stats/_qmc.py:1761-1828, 1862-1917``)
and every function below is asserted bit-equal to the installed SciPy in
``tests/test_oracle_sobol.py``.
"""
from __future__ import annotations

import os

import numpy as np

BITS = 30          # scipy default (``_qmc.py:1780-1781``)
MAXDIM = 21201     # Joe-Kuo table size (``scipy/stats/_sobol.pyi``)
MAXDEG = 18


def load_direction_numbers():
    """Joe-Kuo primitive polynomials and initial direction numbers, read from the data
    file SciPy ships (``scipy/stats/_sobol_direction_numbers.npz``)."""
    import scipy.stats
    path = os.path.join(os.path.dirname(scipy.stats.__file__), "_sobol_direction_numbers.npz")
    with np.load(path) as z:
        return z["poly"].astype(np.int64), z["vinit"].astype(np.int64)


def direction_numbers(d: int, bits: int = BITS) -> np.ndarray:
    """Unscrambled direction numbers ``v[d][bits]`` (uint32), each already shifted so that
    bit ``bits-1`` is the most significant digit (SciPy ``_initialize_v``)."""
    if not 1 <= d <= MAXDIM:
        raise ValueError(f"Maximum supported dimensionality is {MAXDIM}.")
    poly, vinit = load_direction_numbers()
    v = np.zeros((d, bits), dtype=np.uint64)
    v[0, :] = 1
    for k in range(1, d):
        p = int(poly[k])
        deg = p.bit_length() - 1
        for j in range(min(deg, bits)):
            v[k, j] = vinit[k, j]
        for j in range(deg, bits):
            nv = int(v[k, j - deg])
            pow2 = 1
            for q in range(deg):
                pow2 <<= 1
                if (p >> (deg - 1 - q)) & 1:
                    nv ^= pow2 * int(v[k, j - q - 1])
            v[k, j] = nv
    for j in range(bits):
        v[:, j] <<= np.uint64(bits - 1 - j)
    return v.astype(np.uint32)


def scramble_bits(d: int, seed, bits: int = BITS):
    """The random bits SciPy draws for the scramble, in SciPy's order
    (``_qmc.py:1812-1824``): first a ``(d, bits)`` block for the digital shift, then a
    ``(d, bits, bits)`` block for the lower-triangular matrices."""
    rng = np.random.default_rng(seed)
    shift_bits = rng.integers(2, size=(d, bits), dtype=np.uint32)
    ltm_bits = rng.integers(2, size=(d, bits, bits), dtype=np.uint32)
    return shift_bits, ltm_bits


def lms_scramble(v: np.ndarray, ltm_bits: np.ndarray, bits: int = BITS) -> np.ndarray:
    """Left linear matrix scramble of the direction numbers (SciPy ``_cscramble``):
    lower-triangular matrix with unit diagonal over GF(2), applied to every column."""
    d = v.shape[0]
    ltm = np.tril(ltm_bits).astype(np.uint64)
    idx = np.arange(bits)
    ltm[:, idx, idx] = 1
    # row word R_p: column q sits at bit (bits-1-q)
    weights = (np.uint64(1) << (np.uint64(bits - 1) - idx.astype(np.uint64)))
    rows = (ltm * weights[None, None, :]).sum(axis=2).astype(np.uint64)   # (d, bits)
    out = np.zeros((d, bits), dtype=np.uint64)
    v64 = v.astype(np.uint64)
    for p in range(bits):
        masked = rows[:, p][:, None] & v64                    # (d, bits)
        par = np.zeros_like(masked)
        tmp = masked.copy()
        for _ in range(bits):
            par ^= tmp & np.uint64(1)
            tmp >>= np.uint64(1)
        out |= par << np.uint64(bits - 1 - p)
    return out.astype(np.uint32)


def sobol_setup(d: int, seed, bits: int = BITS):
    """``(sv[d][bits], shift[d])`` uint32, equal to SciPy's private ``_sv`` / ``_shift``."""
    v = direction_numbers(d, bits)
    shift_bits, ltm_bits = scramble_bits(d, seed, bits)
    shift = (shift_bits.astype(np.uint64) << np.arange(bits, dtype=np.uint64)[None, :]).sum(axis=1)
    return lms_scramble(v, ltm_bits, bits), shift.astype(np.uint32)


def points_u32(sv: np.ndarray, shift: np.ndarray, start: int, count: int) -> np.ndarray:
    """Integer points ``x_i`` for ``i in [start, start+count)`` by the closed form
    ``x_i = shift ^ XOR_{b in gray(i)} sv[:, b]`` (equivalent to SciPy's ``_draw`` /
    ``fast_forward``, ``_qmc.py:1862-1878, 1932-1957``).  Returns ``(count, d)`` uint32."""
    i = np.arange(start, start + count, dtype=np.uint64)
    gray = i ^ (i >> np.uint64(1))
    x = np.broadcast_to(shift[None, :], (count, sv.shape[0])).copy()
    for b in range(sv.shape[1]):
        on = ((gray >> np.uint64(b)) & np.uint64(1)).astype(bool)
        if on.any():
            x[on] ^= sv[None, :, b]
    return x


# --- the reference's own call form --------------------------------------------------------

def scipy_engine(d: int, seed):
    """The engine exactly as the reference constructs it (``Replicating_Portfolio.py:55``)."""
    from scipy.stats import qmc
    return qmc.Sobol(d, scramble=True, seed=seed)


def scipy_uniform(d: int, seed, start: int, count: int) -> np.ndarray:
    """``(count, d)`` float64 uniforms for the index window, via ``fast_forward`` +
    ``random`` (chunk-continuation is equal to one-shot ``random_base2``)."""
    eng = scipy_engine(d, seed)
    if start:
        eng.fast_forward(start)
    return eng.random(count)


def scipy_points_u32(d: int, seed, start: int, count: int) -> np.ndarray:
    u = scipy_uniform(d, seed, start, count)
    return (u * float(2 ** BITS)).astype(np.uint32)


def sobol_norm(m: int, d: int = 1, seed=1234) -> np.ndarray:
    """``sobol_norm`` as the reference defines it (``Replicating_Portfolio.py:54-57``)."""
    import scipy.stats as stats
    return stats.norm.ppf(scipy_engine(d, seed).random_base2(m))


def normals_window(d: int, seed, start: int, count: int) -> np.ndarray:
    """Reference normals for an index window (used to chunk sizes the reference cannot
    materialise)."""
    import scipy.stats as stats
    return stats.norm.ppf(scipy_uniform(d, seed, start, count))
