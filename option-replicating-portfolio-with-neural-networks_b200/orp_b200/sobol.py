"""Scrambled Sobol engine, bit-exact with ``scipy.stats.qmc.Sobol(d, scramble=True, seed=seed)``
as the reference constructs it in ``sobol_norm`` (``Replicating_Portfolio.py:54-57``).

Host side (here): Joe-Kuo data file + the NumPy ``default_rng(seed)`` bit stream are handed to the
integer set-up routines of ``liborp_b200.so``.  Device side: Gray-code skip-ahead kernels.
"""
from __future__ import annotations

import os
from functools import lru_cache

import numpy as np

from . import _lib

BITS = 30
MAXDIM = 21201


@lru_cache(maxsize=1)
def _joe_kuo():
    """Primitive polynomials / initial direction numbers; SciPy ships the Joe-Kuo table as
    ``scipy/stats/_sobol_direction_numbers.npz`` (data only — no SciPy code runs on this path)."""
    import scipy.stats
    path = os.path.join(os.path.dirname(scipy.stats.__file__), "_sobol_direction_numbers.npz")
    with np.load(path) as z:
        return (np.ascontiguousarray(z["poly"], dtype=np.int64),
                np.ascontiguousarray(z["vinit"], dtype=np.int64))


class SobolEngine:
    """Direction numbers and scramble of one engine; ``d`` must be the reference's
    ``n_time_steps`` because SciPy's scramble of dimension k depends on the total ``d``."""

    def __init__(self, d: int, seed=None, scramble: bool = True):
        d = int(d)
        if d < 1 or d > MAXDIM:
            raise ValueError(f"Maximum supported dimensionality is {MAXDIM}.")
        self.d, self.seed, self.bits = d, seed, BITS
        L = _lib.lib()
        poly, vinit = _joe_kuo()
        v = np.zeros((d, BITS), dtype=np.uint32)
        _lib.check(L.orp_sobol_direction_numbers(_lib.ptr(poly), _lib.ptr(vinit), d, BITS, _lib.ptr(v)))
        if scramble:
            # same calls, same order as qmc.Sobol._scramble (scipy/stats/_qmc.py:1812-1824)
            rng = np.random.default_rng(seed)
            shift_bits = np.ascontiguousarray(rng.integers(2, size=(d, BITS), dtype=np.uint32))
            ltm_bits = np.ascontiguousarray(rng.integers(2, size=(d, BITS, BITS), dtype=np.uint32))
            self.sv = np.zeros((d, BITS), dtype=np.uint32)
            self.shift = np.zeros(d, dtype=np.uint32)
            _lib.check(L.orp_sobol_scramble(d, BITS, _lib.ptr(v), _lib.ptr(shift_bits), _lib.ptr(ltm_bits),
                                            _lib.ptr(self.sv), _lib.ptr(self.shift)))
        else:
            self.sv, self.shift = v, np.zeros(d, dtype=np.uint32)
        self.sv32 = np.zeros((32, (d + 31) & ~31), dtype=np.uint32)     # word-major, dimensions padded to 32
        self.tlane = np.zeros((d, 32, 2), dtype=np.uint32)
        _lib.check(L.orp_sobol_pack(d, BITS, _lib.ptr(self.sv), _lib.ptr(self.shift),
                                    _lib.ptr(self.sv32), _lib.ptr(self.tlane)))
        self._dev = {}

    def device_tables(self, device):
        """``(sv32, tlane)`` int32-typed CUDA tensors holding the uint32 words (cached per device)."""
        torch = _lib.require_cuda()
        device = torch.device(device)
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
        if key not in self._dev:
            self._dev[key] = (torch.from_numpy(self.sv32.view(np.int32)).to(device, non_blocking=False),
                              torch.from_numpy(self.tlane.view(np.int32)).to(device, non_blocking=False))
        return self._dev[key]


@lru_cache(maxsize=16)
def engine(d: int, seed) -> SobolEngine:
    return SobolEngine(d, seed)


def _check_range(start: int, count: int):
    if start < 0 or count < 0:
        raise ValueError("start and count must be >= 0")
    if start + count > 2 ** BITS:
        raise ValueError(f"At most 2**{BITS}={2 ** BITS} distinct points can be generated.")


def sobol_uint32(m: int | None, d: int, seed=None, start: int = 0, count: int | None = None,
                 device="cuda", layout: str = "points"):
    """Integer Sobol points ``x_i`` (``u_i = x_i / 2**30``) for ``i in [start, start+count)``;
    ``count`` defaults to ``2**m - start`` (``random_base2(m)`` when ``start == 0``).
    Returns a ``torch.uint32`` tensor, ``(count, d)`` for ``layout='points'`` (SciPy's) or
    ``(d, count)`` for ``layout='dims'``."""
    torch = _lib.require_cuda()
    if count is None:
        if m is None:
            raise ValueError("give m or count")
        count = 2 ** int(m) - start
    _check_range(start, count)
    eng = engine(int(d), seed)
    device = torch.device(device)
    with torch.cuda.device(device):
        sv32, tlane = eng.device_tables(device)
        shape = (count, eng.d) if layout == "points" else (eng.d, count)
        out = torch.empty(shape, dtype=torch.int32, device=device)
        _lib.check(_lib.lib().orp_sobol_points(_lib.ptr(sv32), _lib.ptr(tlane), eng.d, start, count,
                                               0 if layout == "points" else 1, _lib.ptr(out), _lib.stream_ptr()))
    return out.view(torch.uint32)


def ndtri_u32(x):
    """``ndtri(x / 2**30)`` (fp64) of a CUDA uint32/int32 tensor of integer points."""
    torch = _lib.require_cuda()
    x = x.contiguous()
    with torch.cuda.device(x.device):
        z = torch.empty(x.shape, dtype=torch.float64, device=x.device)
        _lib.check(_lib.lib().orp_ndtri_u32(_lib.ptr(x), x.numel(), _lib.ptr(z), _lib.stream_ptr()))
    return z


def sobol_normal(m: int | None, d: int, seed=None, start: int = 0, count: int | None = None, device="cuda"):
    """``stats.norm.ppf(Sobol(d, scramble=True, seed=seed).random_base2(m))`` — the reference's
    ``sobol_norm`` (``Replicating_Portfolio.py:54-57``) — as a ``(count, d)`` fp64 CUDA tensor."""
    return ndtri_u32(sobol_uint32(m, d, seed, start, count, device))
