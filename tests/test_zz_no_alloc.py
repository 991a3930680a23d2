"""GPU test of ``orp_payoff_reduce_noalloc`` — the entry point for callers that cannot allocate device memory.

Some users of this library price from inside a process that has already given the whole GPU to fixed arenas (or
to another framework's allocator) and is not allowed to call ``cudaMalloc`` any more; they can hand over the path
array and the output slots, but not the scratch buffer ``orp_payoff_reduce`` wants.  The variant tested here keeps
its per-block partial sums in a fixed array inside the library image, limited to
``ORP_PAYOFF_NOALLOC_MAX_OPTIONS`` (8) options per call.

What is checked, through the C ABI: on two calls in a row (different option sets, so the second call finds the
first call's partials in the shared array) the sums, out-of-the-money counts and terminal values equal those of
the regular entry point — to 1e-9 relative, the tolerance the other payoff tests use (they are in fact
bit-identical: same kernels, same fixed-order reduction) — and an over-sized request is refused with a message.

This file sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_OPTIONS = 8


def _both(tm, strikes, cols, kind, s0, with_values):
    """Run the regular and the allocation-free entry point on the same inputs; returns two (sums, otm, values)."""
    from orp_b200 import _lib, payoff_sums
    n_cols, count = tm.shape
    dev = tm.device
    v_ref = torch.zeros(count, dtype=torch.float32, device=dev) if with_values else None
    s_ref, o_ref = payoff_sums(tm, strikes, cols, kind, s0=s0, values_out=v_ref, value_scale=0.01)

    k = np.ascontiguousarray(strikes, dtype=np.float64)
    c = np.ascontiguousarray(cols, dtype=np.int32)
    sums = torch.full((len(k),), float("nan"), dtype=torch.float64, device=dev)
    otm = torch.full((len(k),), float("nan"), dtype=torch.float64, device=dev)
    v_new = torch.zeros(count, dtype=torch.float32, device=dev) if with_values else None
    dtype = _lib.F64 if tm.dtype == torch.float64 else _lib.F32
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().orp_payoff_reduce_noalloc(
            _lib.ptr(tm), dtype, count, count, _lib.PAYOFF[kind], len(k), _lib.ptr(k), _lib.ptr(c), int(n_cols),
            float(s0), 0.01, _lib.ptr(v_new), _lib.ptr(sums), _lib.ptr(otm), _lib.stream_ptr()))
    torch.cuda.synchronize(dev)
    return (s_ref, o_ref, v_ref), (sums, otm, v_new)


def test_noalloc_matches_regular_entry_point_twice_in_a_row(cuda):
    import orp_b200
    p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 12)
    tm64 = orp_b200.simulate_paths("gbm_log", p, m=15, steps=12, device=cuda, time_major=True)
    tm32 = orp_b200.simulate_paths("gbm_log", p, start=5, count=70_001, steps=12, dtype="float32", device=cuda, time_major=True)
    calls = [
        # first call: 8 options (the maximum), calls, fp64 paths, terminal values written
        (tm64, [80.0, 90.0, 95.0, 100.0, 100.0, 105.0, 110.0, 120.0], [12, 12, 6, 12, 3, 12, 9, 12], "call", 100.0, True),
        # second call straight after: fewer options, another payoff, fp32 paths, ragged count
        (tm32, [100.0, 97.5, 130.0], [12, 4, 12], "put", 100.0, False),
        # and once more with the FLOOR payoff of Replicating_Portfolio.py:88-89
        (tm64, [100.0], [12], "floor", 100.0, True),
    ]
    for n_call, (tm, strikes, cols, kind, s0, with_values) in enumerate(calls):
        free_before = torch.cuda.mem_get_info(cuda)[0]
        ref, new = _both(tm, strikes, cols, kind, s0, with_values)
        if n_call > 0:
            # from the second call on nothing may come out of the device heap: the free byte count the driver reports
            # can only move by what THIS test allocates through torch's caching allocator (its outputs: re-used blocks)
            assert torch.cuda.mem_get_info(cuda)[0] >= free_before - (4 << 20)
        np.testing.assert_allclose(new[0].cpu().numpy(), ref[0].cpu().numpy(), rtol=1e-9)
        assert np.array_equal(new[1].cpu().numpy(), ref[1].cpu().numpy())
        if with_values:
            np.testing.assert_allclose(new[2].cpu().numpy(), ref[2].cpu().numpy(), rtol=1e-6)


def test_noalloc_rejects_more_options_than_its_scratch_holds(cuda):
    from orp_b200 import _lib
    import orp_b200
    tm = orp_b200.simulate_paths("gbm_log", dict(S0=100.0, mu=0.05, sigma=0.2, dt=0.25), m=10, steps=4,
                                 device=cuda, time_major=True)
    n = MAX_OPTIONS + 1
    k = np.full(n, 100.0)
    c = np.full(n, 4, dtype=np.int32)
    out = torch.zeros(2 * n, dtype=torch.float64, device=cuda)
    with pytest.raises(ValueError, match="at most 8 options"):
        _lib.check(_lib.lib().orp_payoff_reduce_noalloc(_lib.ptr(tm), _lib.F64, tm.shape[1], tm.shape[1], 0, n, _lib.ptr(k),
                                                        _lib.ptr(c), 5, 100.0, 1.0, None, _lib.ptr(out), _lib.ptr(out[n:]),
                                                        _lib.stream_ptr()))


def test_noalloc_rejects_a_maturity_column_outside_the_path_array(cuda):
    """A bad maturity index must come back as ValueError (the reference's ``paths[:, t]`` would raise IndexError), not
    as an out-of-bounds read on the device; the same check guards the regular entry point."""
    from orp_b200 import _lib
    import orp_b200
    tm = orp_b200.simulate_paths("gbm_log", dict(S0=100.0, mu=0.05, sigma=0.2, dt=0.25), m=10, steps=4,
                                 device=cuda, time_major=True)
    k = np.array([100.0, 100.0])
    out = torch.zeros(4, dtype=torch.float64, device=cuda)
    scratch = torch.zeros(_lib.lib().orp_payoff_scratch_bytes(2), dtype=torch.uint8, device=cuda)
    for bad in ([4, 5], [-1, 4]):
        c = np.array(bad, dtype=np.int32)
        with pytest.raises(ValueError, match="outside"):
            _lib.check(_lib.lib().orp_payoff_reduce_noalloc(_lib.ptr(tm), _lib.F64, tm.shape[1], tm.shape[1], 0, 2, _lib.ptr(k),
                                                            _lib.ptr(c), 5, 100.0, 1.0, None, _lib.ptr(out), _lib.ptr(out[2:]),
                                                            _lib.stream_ptr()))
        with pytest.raises(ValueError, match="outside"):
            _lib.check(_lib.lib().orp_payoff_reduce(_lib.ptr(tm), _lib.F64, tm.shape[1], tm.shape[1], 0, 2, _lib.ptr(k),
                                                    _lib.ptr(c), 5, 100.0, 1.0, None, _lib.ptr(out), _lib.ptr(out[2:]),
                                                    _lib.ptr(scratch), _lib.stream_ptr()))
