"""Per-step risk metrics on the device: the quantiles of the replication residual that the
reference prints as ``VaR: ... (98%), ... (99%)`` at every time step (``Replicating_Portfolio.py:120-122``,
``np.quantile`` over one value per path).  The reference appends every residual to a Python list, which cannot
scale; here the array never leaves the GPU: a radix select over its order-preserving integer keys (four 8-bit
passes pin the k-th smallest element exactly, one more finds its successor for NumPy's linear interpolation).
With ``torch.distributed`` initialised the ranks' histograms are summed between passes, so every rank returns the
quantile of the WHOLE path set.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .dist import allreduce_sum_, is_distributed

BINS = 257


def _key_to_float(key: int) -> float:
    bits = key ^ 0x80000000 if key & 0x80000000 else (~key) & 0xFFFFFFFF
    return float(np.array([bits], dtype=np.uint32).view(np.float32)[0])


def _kth_key(x, k: int, group, ignore_nan: bool = False):
    """Order-preserving key of the k-th smallest element (0-based) over all ranks; None when a NaN is present
    (unless NaNs are ignored: they are never histogrammed, so they simply do not take part)."""
    torch = _lib.require_cuda()
    L = _lib.lib()
    hist = torch.zeros(BINS, dtype=torch.int64, device=x.device)
    prefix = mask = 0
    for shift in (24, 16, 8, 0):
        hist.zero_()
        _lib.check(L.orp_select_hist(_lib.ptr(x), x.numel(), shift, prefix, mask, _lib.ptr(hist), _lib.stream_ptr()))
        allreduce_sum_(hist, group)
        h = hist.cpu().numpy()
        if h[256] > 0 and not ignore_nan:
            return None
        cum = np.cumsum(h[:256])
        d = int(np.searchsorted(cum, k, side="right"))
        if d > 255:
            raise ValueError("rank k is beyond the number of elements")
        k -= int(cum[d - 1]) if d else 0
        prefix |= d << shift
        mask |= 0xFF << shift
    return prefix


def _successor(x, key: int, group):
    """(#elements <= key, smallest key above it) over all ranks."""
    torch = _lib.require_cuda()
    out = torch.tensor([0, 0xFFFFFFFF], dtype=torch.int64, device=x.device)
    _lib.check(_lib.lib().orp_select_next(_lib.ptr(x), x.numel(), key, _lib.ptr(out), _lib.stream_ptr()))
    if is_distributed(group):
        import torch.distributed as dist
        cnt, mn = out[:1].clone(), out[1:].clone()
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
        return int(cnt), int(mn)
    o = out.cpu().numpy()
    return int(o[0]), int(o[1])


def quantiles(x, qs, group=None, ignore_nan: bool = False) -> np.ndarray:
    """``np.quantile(x, qs)`` (linear interpolation) of a contiguous CUDA fp32 vector ``x`` — of the concatenation
    of every rank's ``x`` when distributed.  Returns float64; NaN if any element is NaN, like NumPy — or, with
    ``ignore_nan=True``, ``np.nanquantile``."""
    torch = _lib.require_cuda()
    if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 1:
        raise ValueError("x must be a contiguous 1-D float32 CUDA tensor")
    qs = np.atleast_1d(np.asarray(qs, dtype=np.float64))
    if ((qs < 0) | (qs > 1)).any():
        raise ValueError("Quantiles must be in the range [0, 1]")          # numpy's message
    n = torch.tensor([x.numel()], dtype=torch.int64, device=x.device)
    if ignore_nan:
        n -= torch.isnan(x).sum()
    allreduce_sum_(n, group)
    n = int(n)
    if n == 0:
        return np.full(qs.shape, np.nan)
    out = np.empty(qs.shape)
    with torch.cuda.device(x.device):
        for j, q in enumerate(qs):
            h = (n - 1) * q
            k = int(np.floor(h))
            g = h - k
            key = _kth_key(x, k, group, ignore_nan)
            if key is None:
                out[:] = np.nan
                break
            a = _key_to_float(key)
            b = a
            if k + 1 < n:                            # NumPy forms (b - a) * g even when g == 0 (matters only for infinities)
                le, nxt = _successor(x, key, group)
                if le <= k + 1:                      # no duplicate of a at rank k + 1: the successor is the next distinct value
                    b = _key_to_float(nxt)
            out[j] = a + (b - a) * g if g < 0.5 else b - (b - a) * (1.0 - g)     # numpy's _lerp
    return out


def survivors(lambda_tm, n0: float, dt: float, keep_every: int = 1, seed: int = 1234, start: int = 0):
    """Per-path binomial survivors ``N_t / N`` on the device (``Replicating_Portfolio.py:78-84``): ``lambda_tm`` is the
    TIME-MAJOR ``(n_steps + 1, count)`` mortality-intensity array (``simulate_paths('linear', ...)``); returns the fp32
    ``(n_steps // keep_every + 1, count)`` array of ``N_t / n0`` at the kept columns.  Counter-based draws
    (Philox4x32-10 indexed by global path ``start + j`` and step): independent of the sharding, distributional — not
    bit-wise — parity with the reference's NumPy legacy stream."""
    torch = _lib.require_cuda()
    if lambda_tm.dim() != 2 or not lambda_tm.is_contiguous():
        raise ValueError("lambda_tm must be the contiguous time-major (n_steps + 1, count) tensor")
    if lambda_tm.dtype not in (torch.float64, torch.float32):
        raise ValueError("lambda_tm must be float64 or float32")
    n_cols, count = lambda_tm.shape
    n_steps = n_cols - 1
    if n_steps < 1 or keep_every < 1:
        raise ValueError("need at least one step and keep_every >= 1")
    out = torch.empty((n_steps // keep_every + 1, count), dtype=torch.float32, device=lambda_tm.device)
    with torch.cuda.device(lambda_tm.device):
        _lib.check(_lib.lib().orp_survivors(_lib.ptr(lambda_tm), _lib.F64 if lambda_tm.dtype == torch.float64 else _lib.F32,
                                            count, int(start), count, n_steps, int(keep_every), float(n0), float(dt),
                                            int(seed), _lib.ptr(out), count, _lib.stream_ptr()))
    return out
