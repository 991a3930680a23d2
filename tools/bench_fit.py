#!/usr/bin/env python
"""Kernel-level timing of the fit kernels (CUDA events): samples/s of fit_grad (fwd + bwd) and fit_predict.

    python tools/bench_fit.py [m] [net]        # default 2^25 paths, EO net
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
import orp_b200  # noqa: E402
from orp_b200 import _lib  # noqa: E402
from orp_b200.fit import ReplicationNet, Optimizer, init_params  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 25
net = sys.argv[2] if len(sys.argv) > 2 else "eo"
D = 4
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / 252)
for dtype in ("float64", "float32", "float32 pre-scaled"):
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, dtype=dtype.split()[0], time_major=True)
    inv_scale = 0.01
    if dtype.endswith("pre-scaled"):               # what fit_replication hands to the kernels: fp32 columns already divided by S0
        tm, inv_scale = (tm.double() * 0.01).float(), 1.0
    n = 2 ** m
    vals = torch.empty(n, dtype=torch.float32, device="cuda")
    orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
    f1 = torch.ones(n, dtype=torch.float32, device="cuda") if net == "rp" else None
    f2 = torch.zeros(n, dtype=torch.float32, device="cuda") if net == "rp" else None
    model = ReplicationNet(net, init_params(net, 1234, 0.1 if net == "eo" else [0.9, 0.1]), "cuda")
    opt = Optimizer("cuda")
    b = model.batch(tm[D - 1], tm[D], vals, inv_scale, 1.0, 1.0, feat1=f1, feat2=f2)
    L = _lib.lib()
    st = _lib.stream_ptr()
    _lib.check(L.orp_fit_begin(_lib.ptr(opt.state), 10 ** 6, 10 ** 6, 0, st))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for parts in (1, 8):
        chunks = model.split(b, parts)
        for rep in range(2):
            e0.record()
            for _ in range(10):
                for c in chunks:
                    _lib.check(L.orp_fit_grad(C.byref(c), _lib.ptr(model.params), _lib.ptr(opt.state), _lib.ptr(model.gsum),
                                              _lib.ptr(model.scratch), st))
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"{net} {dtype} fit_grad   parts={parts}  {ms:8.3f} ms/pass  {n / ms / 1e6:8.2f} G samples/s", flush=True)
    out = torch.empty_like(vals)
    for rep in range(2):
        e0.record()
        for _ in range(10):
            _lib.check(L.orp_fit_predict(C.byref(b), _lib.ptr(model.params), _lib.ptr(out), _lib.ptr(model.sums),
                                         _lib.ptr(model.scratch), st))
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    bytes_ = n * (2 * tm.element_size() + 4 + 4)
    print(f"{net} {dtype} fit_predict           {ms:8.3f} ms/pass  {n / ms / 1e6:8.2f} G samples/s  {bytes_ / ms / 1e6:8.1f} GB/s", flush=True)
    del tm
