"""ctypes binding of ``liborp_b200.so`` (C ABI declared in ``include/orp_b200.h``).

There is no CPU fallback: if the shared library is missing and cannot be built, importing the
compute entry points raises.  torch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("ORP_B200_LIB") or os.path.join(_PKG, "liborp_b200.so")      # override: instrumented builds

ORP_OK, ORP_E_VALUE, ORP_E_CUDA, ORP_E_STATE = 0, -1, -2, -3
MODEL = {"gbm_log": 0, "gbm_arith": 1, "gbm_euler": 1, "sv": 2, "linear": 3}
ORDER = {"incremental": 0, "bridge": 1}
PAYOFF = {"call": 0, "put": 1, "floor": 2}
NET = {"eo": 0, "rp": 1}
LOSS = {"mse": 0, "pinball": 1}
F64, F32 = 0, 1
MAX_OPTIONS = 64
FIT_MAX_PARAMS = 128
FIT_STATE_DOUBLES = 1024
FIT_MAX_SEGMENTS = 8
GSUM_LEN = FIT_MAX_PARAMS + 2

# indices into the device-resident fit state (mirror of the enum in csrc/orp_fit.cu)
ST_LR, ST_ADAM_T, ST_EPOCH, ST_EPOCHS, ST_PATIENCE, ST_USE_SCHED, ST_WAIT, ST_BEST, ST_STOP, \
    ST_STOPPED_EARLY, ST_LAST_LOSS, ST_EPOCHS_RUN, ST_HAVE_BEST, ST_TOTAL_EPOCHS = range(14)
ST_ADAM_M = 16
ST_ADAM_V = ST_ADAM_M + FIT_MAX_PARAMS
ST_BEST_W = ST_ADAM_V + FIT_MAX_PARAMS
ST_TICKET = ST_BEST_W + FIT_MAX_PARAMS
ST_ABORT = ST_TICKET + 2


class PathParams(C.Structure):
    _fields_ = [("model", C.c_int32), ("ordering", C.c_int32), ("n_steps", C.c_int32), ("keep_every", C.c_int32),
                ("out_dtype", C.c_int32), ("sobol_dim", C.c_int32),
                ("s0", C.c_double), ("mu", C.c_double), ("sigma", C.c_double), ("dt", C.c_double),
                ("v0", C.c_double), ("a", C.c_double), ("b", C.c_double), ("c", C.c_double)]


class FitBatch(C.Structure):
    _fields_ = [("net", C.c_int32), ("loss", C.c_int32), ("paths_dtype", C.c_int32), ("reserved", C.c_int32),
                ("count", C.c_uint64), ("d_s_t", C.c_void_p), ("d_s_t1", C.c_void_p),
                ("d_feat1", C.c_void_p), ("d_feat2", C.c_void_p), ("d_target", C.c_void_p),
                ("inv_scale", C.c_double), ("bond_t", C.c_double), ("bond_t1", C.c_double),
                ("n_seg", C.c_uint32), ("reserved2", C.c_uint32), ("seg_stride", C.c_uint64)]


_vp, _u64, _i32, _f64 = C.c_void_p, C.c_uint64, C.c_int, C.c_double

# name -> (restype, argtypes); kept in step with include/orp_b200.h (tests/test_abi.py checks both ways)
PROTOTYPES = {
    "orp_version": (C.c_int, []),
    "orp_last_error": (C.c_char_p, []),
    "orp_sobol_direction_numbers": (C.c_int, [_vp, _vp, _i32, _i32, _vp]),
    "orp_sobol_scramble": (C.c_int, [_i32, _i32, _vp, _vp, _vp, _vp, _vp]),
    "orp_sobol_pack": (C.c_int, [_i32, _i32, _vp, _vp, _vp, _vp]),
    "orp_sobol_points": (C.c_int, [_vp, _vp, _i32, _u64, _u64, _i32, _vp, _vp]),
    "orp_ndtri_u32": (C.c_int, [_vp, _u64, _vp, _vp]),
    "orp_paths": (C.c_int, [C.POINTER(PathParams), _vp, _vp, _vp, _vp, _vp, _u64, _u64, _vp, _u64, _vp]),
    "orp_bridge_schedule_bytes": (C.c_uint64, [_i32]),
    "orp_bridge_schedule": (C.c_int, [_i32, _vp]),
    "orp_payoff_scratch_bytes": (C.c_uint64, [_i32]),
    "orp_payoff_reduce": (C.c_int, [_vp, _i32, _u64, _u64, _i32, _i32, _vp, _vp, _i32, _f64, _f64, _vp, _vp, _vp, _vp, _vp]),
    "orp_payoff_reduce_noalloc": (C.c_int, [_vp, _i32, _u64, _u64, _i32, _i32, _vp, _vp, _i32, _f64, _f64, _vp, _vp, _vp, _vp]),
    "orp_paths_payoff_scratch_bytes": (C.c_uint64, [_i32]),
    "orp_paths_payoff": (C.c_int, [C.POINTER(PathParams), _vp, _vp, _vp, _vp, _u64, _u64, _i32, _i32, _vp, _vp, _f64, _vp, _vp, _vp, _vp]),
    "orp_fit_n_params": (C.c_int, [_i32]),
    "orp_fit_scratch_bytes": (C.c_uint64, []),
    "orp_fit_grad": (C.c_int, [C.POINTER(FitBatch), _vp, _vp, _vp, _vp, _vp]),
    "orp_fit_update": (C.c_int, [_i32, _vp, _vp, _vp, _i32, _vp]),
    "orp_fit_step": (C.c_int, [C.POINTER(FitBatch), _vp, _vp, _vp, _vp, _i32, _vp]),
    "orp_fit_epochs": (C.c_int, [C.POINTER(FitBatch), _i32, _vp, _vp, _vp, _vp, _i32, _vp]),
    "orp_fit_run": (C.c_int, [C.POINTER(FitBatch), _i32, _vp, _vp, _vp, _vp, _i32, _vp]),
    "orp_fit_run_peers": (C.c_int, [C.POINTER(FitBatch), _i32, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "orp_peer_buffer_bytes": (C.c_uint64, []),
    "orp_peer_buffer_create": (C.c_int, [C.POINTER(C.c_void_p), _vp]),
    "orp_peer_buffer_open": (C.c_int, [_vp, C.POINTER(C.c_void_p)]),
    "orp_peer_buffer_close": (C.c_int, [_vp]),
    "orp_peer_buffer_destroy": (C.c_int, [_vp]),
    "orp_fit_scale_column": (C.c_int, [_vp, _i32, _u64, _f64, _vp, _vp]),
    "orp_fit_begin": (C.c_int, [_vp, _i32, _i32, _i32, _vp]),
    "orp_fit_init": (C.c_int, [_vp, _vp]),
    "orp_fit_end": (C.c_int, [_i32, _vp, _vp, _vp]),
    "orp_fit_predict": (C.c_int, [C.POINTER(FitBatch), _vp, _vp, _vp, _vp, _vp]),
    "orp_fit_residuals": (C.c_int, [C.POINTER(FitBatch), _vp, _vp, _vp]),
    "orp_select_hist": (C.c_int, [_vp, _u64, _i32, C.c_uint32, C.c_uint32, _vp, _vp]),
    "orp_select_next": (C.c_int, [_vp, _u64, C.c_uint32, _vp, _vp]),
    "orp_select_hist_multi": (C.c_int, [_vp, _u64, _i32, _i32, _vp, _vp, _vp]),
    "orp_select_pick": (C.c_int, [_vp, _i32, _i32, _vp, _vp]),
    "orp_select_next_multi": (C.c_int, [_vp, _u64, _i32, _vp, _vp, _vp]),
    "orp_survivors": (C.c_int, [_vp, _i32, _u64, _u64, _u64, _i32, _i32, _f64, _f64, _u64, _vp, _u64, _vp]),
    "orp_fit_epoch_minibatch": (C.c_int, [C.POINTER(FitBatch), _vp, _vp, _i32, _u64, _vp]),
    "orp_probe_throughput": (C.c_int, [_i32, _i32, _vp, C.POINTER(C.c_double), _vp]),
}

_lib = None


def _build_if_needed() -> None:
    sys.path.insert(0, _PKG)
    try:
        import build as _build          # option-replicating-portfolio-with-neural-networks_b200/build.py
        try:
            _build.build()
        except RuntimeError:
            if not os.path.exists(LIB_PATH):
                raise
    finally:
        sys.path.remove(_PKG)


def lib() -> C.CDLL:
    """The loaded library; built in-tree on first use when nvcc is present."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH) or os.environ.get("ORP_B200_REBUILD"):
            _build_if_needed()
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing and could not be built; there is no CPU fallback")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int) -> None:
    """Map the C error convention onto the reference's (Python exceptions; ValueError for bad
    arguments as SciPy raises, scipy/stats/_qmc.py:1771-1774)."""
    if rc == ORP_OK:
        return
    msg = lib().orp_last_error().decode("utf-8", "replace")
    if rc == ORP_E_VALUE:
        raise ValueError(msg)
    raise RuntimeError(f"liborp_b200 error {rc}: {msg}")


def ptr(t) -> int:
    """Raw pointer of a torch tensor / numpy array (None -> NULL)."""
    if t is None:
        return 0
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("orp_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch
