"""Per-step risk metrics on the device: the quantiles of the replication residual that the
reference prints as ``VaR: ... (98%), ... (99%)`` at every time step (``Replicating_Portfolio.py:120-122``,
``np.quantile`` over one value per path).  The reference appends every residual to a Python list, which cannot
scale; here the array never leaves the GPU: a radix select over its order-preserving integer keys (four 8-bit
passes pin the k-th smallest element exactly, one more finds its successor for NumPy's linear interpolation).
All requested quantiles share the five passes, the digit is picked on the device between passes
(``orp_select_pick``) and the host reads the result once.  With ``torch.distributed`` initialised the ranks'
histograms are summed between passes, so every rank returns the quantile of the WHOLE path set.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .dist import allreduce_sum_, is_distributed

BINS = 257


def _key_to_float(key: int) -> float:
    bits = key ^ 0x80000000 if key & 0x80000000 else (~key) & 0xFFFFFFFF
    return float(np.array([bits], dtype=np.uint32).view(np.float32)[0])


MAX_Q = 4                      # ORP_SELECT_MAX_Q: order statistics per set of passes


def _select(x, ks, group):
    """Order-preserving keys of the ``ks``-th smallest elements (0-based ranks over all ranks' elements; at most
    ``MAX_Q``), with the count of elements <= each key and the smallest key above it: ``(keys, le, nxt, nan, bad)``.
    Nothing is read back before the last pass: the selection state lives in device memory."""
    torch = _lib.require_cuda()
    import torch.distributed as dist
    L = _lib.lib()
    Q = len(ks)
    st_host = np.zeros((Q, 4), dtype=np.uint64)            # orp_select_state: k | prefix, mask | nan, bad | reserved
    st_host[:, 0] = ks
    state = torch.from_numpy(st_host.view(np.int64)).to(x.device)
    hist = torch.zeros(Q * BINS, dtype=torch.int64, device=x.device)
    out = torch.tensor([0, 0xFFFFFFFF] * Q, dtype=torch.int64, device=x.device)
    sp = _lib.stream_ptr()
    for shift in (24, 16, 8, 0):
        _lib.check(L.orp_select_hist_multi(_lib.ptr(x), x.numel(), shift, Q, _lib.ptr(state), _lib.ptr(hist), sp))
        allreduce_sum_(hist, group)
        _lib.check(L.orp_select_pick(_lib.ptr(hist), shift, Q, _lib.ptr(state), sp))
    _lib.check(L.orp_select_next_multi(_lib.ptr(x), x.numel(), Q, _lib.ptr(state), _lib.ptr(out), sp))
    o = out.view(Q, 2)
    if is_distributed(group):
        cnt, mn = o[:, 0].contiguous(), o[:, 1].contiguous()
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=group)
        o = torch.stack([cnt, mn], dim=1)
    both = torch.cat([state.view(-1), o.reshape(-1)]).cpu().numpy()              # the one read-back
    st = both[:4 * Q].view(np.uint64).reshape(Q, 4)
    o = both[4 * Q:].reshape(Q, 2)
    keys = (st[:, 1] & np.uint64(0xFFFFFFFF)).astype(np.int64)
    nan = ((st[:, 2] & np.uint64(0xFFFFFFFF)) != 0)
    bad = ((st[:, 2] >> np.uint64(32)) != 0)
    return keys, o[:, 0], o[:, 1], nan, bad


def quantiles(x, qs, group=None, ignore_nan: bool = False, total: int | None = None) -> np.ndarray:
    """``np.quantile(x, qs)`` (linear interpolation) of a contiguous CUDA fp32 vector ``x`` — of the concatenation
    of every rank's ``x`` when distributed.  Returns float64; NaN if any element is NaN, like NumPy — or, with
    ``ignore_nan=True``, ``np.nanquantile``.  ``total``: the number of elements over all ranks when the caller knows
    it (saves the one collective + read-back that counts them)."""
    torch = _lib.require_cuda()
    if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 1:
        raise ValueError("x must be a contiguous 1-D float32 CUDA tensor")
    qs = np.atleast_1d(np.asarray(qs, dtype=np.float64))
    if ((qs < 0) | (qs > 1)).any():
        raise ValueError("Quantiles must be in the range [0, 1]")          # numpy's message
    if total is not None and not ignore_nan:
        n = int(total)
    elif not ignore_nan and not is_distributed(group):
        n = x.numel()
    else:
        n = torch.tensor([x.numel()], dtype=torch.int64, device=x.device)
        if ignore_nan:
            n -= torch.isnan(x).sum()
        allreduce_sum_(n, group)
        n = int(n)
    if n == 0:
        return np.full(qs.shape, np.nan)
    out = np.empty(qs.shape)
    flat = out.reshape(-1)
    qf = qs.reshape(-1)
    with torch.cuda.device(x.device):
        for j0 in range(0, qf.size, MAX_Q):
            h = (n - 1) * qf[j0:j0 + MAX_Q]
            k = np.floor(h).astype(np.int64)
            g = h - k
            keys, le, nxt, nan, bad = _select(x, k, group)
            if nan.any() and not ignore_nan:           # (first: NaNs are not histogrammed, so they also leave k "beyond")
                out[...] = np.nan
                break
            if bad.any():
                raise ValueError("rank k is beyond the number of elements")
            for j in range(k.size):
                a = _key_to_float(int(keys[j]))
                b = a
                # NumPy forms (b - a) * g even when g == 0 (matters only for infinities); no duplicate of a at rank
                # k + 1 means the successor is the next distinct value
                if k[j] + 1 < n and le[j] <= k[j] + 1:
                    b = _key_to_float(int(nxt[j]))
                flat[j0 + j] = a + (b - a) * g[j] if g[j] < 0.5 else b - (b - a) * (1.0 - g[j])     # numpy's _lerp
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
