"""Path generation, payoff and estimator reductions on the GPU (host-side mirror of the reference's
simulation block, ``Replicating_Portfolio.py:54-96`` / ``:268-293`` / ``European Options.ipynb`` cells 6-8).
"""
from __future__ import annotations

import ctypes as C
from math import ceil, floor

import numpy as np

from . import _lib
from .sobol import engine, _check_range

_DTYPES = {"float64": _lib.F64, "fp64": _lib.F64, "f64": _lib.F64, "float32": _lib.F32, "fp32": _lib.F32, "f32": _lib.F32}


def n_time_steps_of(T: float, dt: float) -> int:
    """``ceil(T/dt)+1`` (``Replicating_Portfolio.py:51``)."""
    return ceil(T / dt) + 1


def reduction_of(n_time_steps: int, T: float, rebalancing: float) -> int:
    """``floor((n_time_steps-1)/(T/rebalancing))`` (``Replicating_Portfolio.py:92``)."""
    return floor((n_time_steps - 1) / (T / rebalancing))


def bond_curve(r: float, T: float, n_time_steps: int, keep_every: int = 1) -> np.ndarray:
    """``exp(r*linspace(0,T,n))[::reduction]`` (``Replicating_Portfolio.py:68, 94``): a per-time scalar,
    never materialised per path."""
    return np.exp(r * np.linspace(0, T, n_time_steps))[::keep_every]


def shard_range(total: int, rank: int, world: int):
    """Contiguous Sobol index range of ``rank``: the union over ranks is exactly the 1-GPU point set."""
    per = -(-total // world)
    start = min(rank * per, total)
    return start, min(per, total - start)


_bridge_cache = {}


def _bridge_schedule(n_steps: int, device):
    torch = _lib.require_cuda()
    key = (n_steps, str(device))
    if key not in _bridge_cache:
        L = _lib.lib()
        buf = np.zeros(L.orp_bridge_schedule_bytes(n_steps), dtype=np.uint8)
        _lib.check(L.orp_bridge_schedule(n_steps, _lib.ptr(buf)))
        _bridge_cache[key] = torch.from_numpy(buf).to(device)
    return _bridge_cache[key]


def bridge_schedule_host(n_steps: int) -> np.ndarray:
    """The depth-first bridge program as a structured array (dim, emits, wl, wr, sd); host only."""
    L = _lib.lib()
    buf = np.zeros(L.orp_bridge_schedule_bytes(n_steps), dtype=np.uint8)
    _lib.check(L.orp_bridge_schedule(n_steps, _lib.ptr(buf)))
    return buf.view(np.dtype([("dim", "<i4"), ("emits", "<i4"), ("wl", "<f8"), ("wr", "<f8"), ("sd", "<f8")]))


def simulate_paths(model: str, params: dict, m: int | None = None, steps: int | None = None, seed: int = 1235,
                   ordering: str = "incremental", dtype: str = "float64", keep_every: int = 1,
                   shard=(0, 1), start: int | None = None, count: int | None = None, seed_sv: int = 1234,
                   device="cuda", out=None, time_major: bool = False, engine_dim: int | None = None):
    """Simulate ``count`` paths of ``steps`` fine steps and keep every ``keep_every``-th column.

    ``model``: ``'gbm_log'`` (EO cell 6), ``'gbm_arith'``/``'gbm_euler'`` (``Replicating_Portfolio.py:65``),
    ``'sv'`` (``:285-287``).  ``params``: ``S0`` (or ``Y``), ``mu`` (or ``r``), ``sigma``, ``dt``; for ``'sv'``
    ``s0, a, b, c``; ``'linear'`` (``:76``) takes ``S0=l0, mu=c, sigma=ita``.  The engines have ``d = steps + 1`` dimensions
    (``engine_dim`` overrides it when fewer steps than the reference's grid are run) and seeds 1235 (price) / 1234
    (volatility) as in the reference.  Index range: ``[start, start+count)`` or the ``shard=(rank, world)``
    slice of ``2**m`` paths.

    Returns a tensor shaped ``(count, n_cols)`` like the reference's ``Y_paths`` — a transposed view of
    the time-major storage (``paths[:, t]`` is contiguous) — or the ``(n_cols, count)`` storage itself
    with ``time_major=True``."""
    torch = _lib.require_cuda()
    if steps is None:
        raise ValueError("steps is required")
    steps = int(steps)
    if start is None:
        if m is None:
            raise ValueError("give m or start/count")
        start, count = shard_range(2 ** int(m), int(shard[0]), int(shard[1]))
    _check_range(start, count)
    if dtype not in _DTYPES:
        raise ValueError(f"dtype must be one of {sorted(_DTYPES)}")
    if model not in _lib.MODEL:
        raise ValueError(f"model must be one of {sorted(_lib.MODEL)}")
    if ordering not in _lib.ORDER:
        raise ValueError(f"ordering must be one of {sorted(_lib.ORDER)}")
    keep_every = int(keep_every)
    if keep_every < 1:
        raise ValueError("keep_every must be >= 1")
    d = steps + 1 if engine_dim is None else int(engine_dim)     # the reference's d = n_time_steps
    if steps > d - 1:
        raise ValueError("steps must be <= engine_dim - 1")
    n_cols = steps // keep_every + 1
    device = torch.device(device)
    p = _path_params(model, params, steps, ordering, dtype, keep_every, d)
    tdt = torch.float64 if p.out_dtype == _lib.F64 else torch.float32
    with torch.cuda.device(device):
        sv_a, tl_a = engine(d, seed).device_tables(device)
        sv_b = tl_b = None
        if model == "sv":
            sv_b, tl_b = engine(d, seed_sv).device_tables(device)
        sched = _bridge_schedule(steps, device) if ordering == "bridge" else None
        if out is None:
            out = torch.empty((n_cols, count), dtype=tdt, device=device)
        elif out.shape != (n_cols, count) or out.dtype != tdt or not out.is_contiguous():
            raise ValueError("out must be a contiguous (n_cols, count) tensor of the requested dtype")
        _lib.check(_lib.lib().orp_paths(C.byref(p), _lib.ptr(sv_a), _lib.ptr(tl_a), _lib.ptr(sv_b), _lib.ptr(tl_b),
                                        _lib.ptr(sched), start, count, _lib.ptr(out), count, _lib.stream_ptr()))
    return out if time_major else out.t()


_scratch = {}
_sweep_plans = {}


def _payoff_scratch(device, n_opt):
    torch = _lib.require_cuda()
    key = str(device)
    need = _lib.lib().orp_payoff_scratch_bytes(_lib.MAX_OPTIONS)
    if key not in _scratch:
        _scratch[key] = torch.empty(need, dtype=torch.uint8, device=device)
    return _scratch[key]


def _sweep_plan(strikes, columns, n_cols, device):
    """Host-side preparation of one (strikes, columns) set, cached: the kernel handles options in chunks of 8 and reads a
    maturity column once per chunk it appears in, so the options are grouped by column (option 0 stays first: it owns
    values_out) — a 64-option sweep over 8 maturities streams 8 columns instead of 64 — and a packed output buffer
    [sums | otm | count] that is all-reduced as one vector.  Results are returned in the caller's order."""
    torch = _lib.require_cuda()
    strikes = np.ascontiguousarray(np.atleast_1d(strikes), dtype=np.float64)
    columns = np.ascontiguousarray(np.atleast_1d(columns), dtype=np.int64)
    key = (strikes.tobytes(), columns.tobytes(), int(n_cols), str(device))
    plan = _sweep_plans.get(key)
    if plan is None:
        if strikes.shape != columns.shape:
            raise ValueError("strikes and columns must have the same length")
        n_opt = strikes.size
        if not 1 <= n_opt <= _lib.MAX_OPTIONS:
            raise ValueError(f"between 1 and {_lib.MAX_OPTIONS} options per call")
        cols = np.where(columns < 0, columns + n_cols, columns).astype(np.int32)
        if (cols < 0).any() or (cols >= n_cols).any():
            raise ValueError("column index out of range")
        order = np.concatenate([[0], 1 + np.argsort(cols[1:], kind="stable")]).astype(np.int64)
        identity = np.array_equal(order, np.arange(n_opt))
        plan = dict(n_opt=n_opt, cols=cols, strikes_s=np.ascontiguousarray(strikes[order]), cols_s=np.ascontiguousarray(cols[order]),
                    inv=None if identity else torch.from_numpy(np.argsort(order)).to(device), inv_host=np.argsort(order),
                    packed=torch.empty(2 * n_opt + 1, dtype=torch.float64, device=device))
        if len(_sweep_plans) > 256:
            _sweep_plans.clear()
        _sweep_plans[key] = plan
    return plan


def _payoff_launch(paths_tm, plan, kind, s0, values_out, value_scale):
    """Enqueues the payoff kernels of a plan: sums -> plan['packed'][:n], otm -> [n:2n], in the plan's (grouped) order."""
    torch = _lib.require_cuda()
    if kind not in _lib.PAYOFF:
        raise ValueError(f"kind must be one of {sorted(_lib.PAYOFF)}")
    if not paths_tm.is_contiguous():
        raise ValueError("paths_tm must be the contiguous time-major tensor")
    n_cols, count = paths_tm.shape
    n_opt, packed = plan["n_opt"], plan["packed"]
    dt = _lib.F64 if paths_tm.dtype == torch.float64 else _lib.F32
    with torch.cuda.device(paths_tm.device):
        _lib.check(_lib.lib().orp_payoff_reduce(
            _lib.ptr(paths_tm), dt, count, count, _lib.PAYOFF[kind], n_opt, _lib.ptr(plan["strikes_s"]), _lib.ptr(plan["cols_s"]),
            int(n_cols), float(s0), float(value_scale), _lib.ptr(values_out), _lib.ptr(packed), _lib.ptr(packed) + 8 * n_opt,
            _lib.ptr(_payoff_scratch(paths_tm.device, n_opt)), _lib.stream_ptr()))


def payoff_sums(paths_tm, strikes, columns, kind: str = "call", s0: float = 0.0,
                values_out=None, value_scale: float = 1.0):
    """Per option ``j``: ``sum_i payoff(S[columns[j]][i], strikes[j])`` and the out-of-the-money count, as
    two fp64 CUDA tensors of length ``n_opt`` (no host sync).  ``paths_tm`` is the TIME-MAJOR
    ``(n_cols, count)`` tensor.  ``values_out`` (fp32, ``count``) receives option 0's payoff times
    ``value_scale`` — ``values[:, -1]`` of the backward loop."""
    plan = _sweep_plan(strikes, columns, paths_tm.shape[0], paths_tm.device)
    _payoff_launch(paths_tm, plan, kind, s0, values_out, value_scale)
    n = plan["n_opt"]
    out = plan["packed"][:2 * n].clone()                    # the plan's buffer is overwritten by the next call
    sums, otm = out[:n], out[n:]
    if plan["inv"] is not None:
        sums, otm = sums[plan["inv"]], otm[plan["inv"]]
    return sums, otm


def price_european(paths, strikes, maturities_idx, kind: str = "call", r: float = 0.0, dt: float = 1.0,
                   total_paths: int | None = None, group=None):
    """Discounted Monte-Carlo prices ``exp(-r * t_col) * mean(payoff)`` for up to 64 (strike, maturity
    column) pairs sharing one path set.  ``paths`` is what :func:`simulate_paths` returned (either
    orientation).  ``dt`` is the spacing of the KEPT columns.  With ``torch.distributed`` initialised (or
    ``group`` given) the packed sums are all-reduced over ranks — ONE collective per call, into a buffer that lives with
    the cached option set, so a repeated sweep allocates nothing.  Returns ``(prices, prob_otm)`` NumPy arrays."""
    tm = paths if paths.is_contiguous() else paths.t()
    plan = _sweep_plan(strikes, maturities_idx, tm.shape[0], tm.device)
    _payoff_launch(tm, plan, kind, 0.0, None, 1.0)
    k, packed = plan["n_opt"], plan["packed"]
    packed[2 * k] = float(tm.shape[1])
    from .dist import allreduce_sum_
    allreduce_sum_(packed, group)
    host = packed.cpu().numpy()
    total = host[-1] if total_paths is None else float(total_paths)
    inv = plan["inv_host"]
    disc = np.exp(-r * dt * plan["cols"])
    return disc * host[:k][inv] / total, host[k:2 * k][inv] / total


def _path_params(model, params, steps, ordering, dtype, keep_every, engine_dim):
    p = _lib.PathParams()
    p.model, p.ordering, p.n_steps, p.keep_every = _lib.MODEL[model], _lib.ORDER[ordering], steps, keep_every
    p.out_dtype = _DTYPES[dtype]
    p.sobol_dim = steps + 1 if engine_dim is None else int(engine_dim)
    p.s0 = float(params.get("S0", params.get("Y", 1.0)))
    p.mu = float(params.get("mu", params.get("r", 0.0)))
    p.sigma = float(params.get("sigma", 0.0))
    p.dt = float(params["dt"])
    p.v0, p.a, p.b, p.c = (float(params.get(k, 0.0)) for k in ("s0", "a", "b", "c"))
    return p


_fused_buffers = {}


def payoff_sums_fused(model: str, params: dict, strikes, maturity_steps, kind: str = "call", m: int | None = None,
                      steps: int | None = None, seed: int = 1235, shard=(0, 1), start: int | None = None,
                      count: int | None = None, seed_sv: int = 1234, device="cuda", s0: float = 0.0,
                      engine_dim: int | None = None):
    """Generate the paths and pay the options off in ONE kernel, without storing a single path
    (``orp_paths_payoff``): per option ``j`` the sum of ``payoff(S(maturity_steps[j]), strikes[j])`` over the index
    range and the out-of-the-money count, as two fp64 CUDA tensors (views of per-device buffers that the next call
    overwrites; no host sync).  Same models, engines and index-range arguments as :func:`simulate_paths`;
    ``maturity_steps`` are fine-grid time steps in ``[0, steps]``.  Up to 64 options per call."""
    torch = _lib.require_cuda()
    if steps is None:
        raise ValueError("steps is required")
    steps = int(steps)
    if start is None:
        if m is None:
            raise ValueError("give m or start/count")
        start, count = shard_range(2 ** int(m), int(shard[0]), int(shard[1]))
    _check_range(start, count)
    if model not in _lib.MODEL:
        raise ValueError(f"model must be one of {sorted(_lib.MODEL)}")
    if kind not in _lib.PAYOFF:
        raise ValueError(f"kind must be one of {sorted(_lib.PAYOFF)}")
    strikes = np.ascontiguousarray(np.atleast_1d(strikes), dtype=np.float64)
    mats = np.ascontiguousarray(np.atleast_1d(maturity_steps), dtype=np.int64)
    if strikes.shape != mats.shape:
        raise ValueError("strikes and maturity_steps must have the same length")
    n_opt = strikes.size
    if not 1 <= n_opt <= _lib.MAX_OPTIONS:
        raise ValueError(f"between 1 and {_lib.MAX_OPTIONS} options per call")
    mats = np.where(mats < 0, mats + steps + 1, mats)
    if (mats < 0).any() or (mats > steps).any():
        raise ValueError("maturity step out of range")
    mats = np.ascontiguousarray(mats, dtype=np.int32)
    d = steps + 1 if engine_dim is None else int(engine_dim)
    if steps > d - 1:
        raise ValueError("steps must be <= engine_dim - 1")
    device = torch.device(device)
    p = _path_params(model, params, steps, "incremental", "float64", 1, d)
    with torch.cuda.device(device):
        sv_a, tl_a = engine(d, seed).device_tables(device)
        sv_b = tl_b = None
        if model == "sv":
            sv_b, tl_b = engine(d, seed_sv).device_tables(device)
        key = str(device)
        if key not in _fused_buffers:          # scratch and outputs live as long as the process: no allocation per call
            _fused_buffers[key] = (torch.empty(_lib.lib().orp_paths_payoff_scratch_bytes(_lib.MAX_OPTIONS), dtype=torch.uint8, device=device),
                                   torch.empty(2 * _lib.MAX_OPTIONS + 1, dtype=torch.float64, device=device))
        scratch, out = _fused_buffers[key]
        _lib.check(_lib.lib().orp_paths_payoff(C.byref(p), _lib.ptr(sv_a), _lib.ptr(tl_a), _lib.ptr(sv_b), _lib.ptr(tl_b),
                                               start, count, _lib.PAYOFF[kind], n_opt, _lib.ptr(strikes), _lib.ptr(mats), float(s0),
                                               _lib.ptr(out), _lib.ptr(out) + 8 * _lib.MAX_OPTIONS, _lib.ptr(scratch), _lib.stream_ptr()))
    return out[:n_opt], out[_lib.MAX_OPTIONS:_lib.MAX_OPTIONS + n_opt]


def price_european_fused(model: str, params: dict, strikes, maturity_steps, kind: str = "call", r: float = 0.0,
                         m: int | None = None, steps: int | None = None, total_paths: int | None = None, group=None, **kw):
    """Discounted Monte-Carlo prices ``exp(-r * dt * step) * mean(payoff)`` of up to 64 (strike, maturity step) pairs
    straight from the Sobol engine: paths generated, paid off and reduced in one kernel, never stored
    (:func:`payoff_sums_fused`); ONE all-reduce of the packed sums when ``torch.distributed`` is initialised.  Pass
    ``shard=(rank, world)`` (or ``start``/``count``) as for :func:`simulate_paths`.  Returns ``(prices, prob_otm)``."""
    torch = _lib.require_cuda()
    sums, otm = payoff_sums_fused(model, params, strikes, maturity_steps, kind, m=m, steps=steps, **kw)
    k = sums.numel()
    packed = sums.new_empty(2 * k + 1)
    packed[:k], packed[k:2 * k] = sums, otm
    if kw.get("start") is not None:
        n_local = kw["count"]
    else:
        n_local = shard_range(2 ** int(m), *(int(x) for x in kw.get("shard", (0, 1))))[1]
    packed[2 * k] = float(n_local)
    from .dist import allreduce_sum_
    allreduce_sum_(packed, group)
    packed = packed.cpu().numpy()
    total = packed[-1] if total_paths is None else float(total_paths)
    mats = np.atleast_1d(np.asarray(maturity_steps))
    mats = np.where(mats < 0, mats + int(steps) + 1, mats)
    disc = np.exp(-r * float(params["dt"]) * mats)
    return disc * packed[:k] / total, packed[k:2 * k] / total


