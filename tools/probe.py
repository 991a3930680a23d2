#!/usr/bin/env python
"""Measured issue rates of the instruction kinds the fit kernel is made of (orp_probe_throughput): warp instructions per
cycle and SM sub-partition at the clock nvidia-smi reports, and TFLOP/s for the FMA kinds.

    python tools/probe.py
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
from orp_b200 import _lib  # noqa: E402


def probe(kind: int, iters: int = 2000, reps: int = 5):
    """Best-of-`reps` warp instructions per second of one probe kind on the current device."""
    L = _lib.lib()
    sink = torch.zeros(4, dtype=torch.float32, device="cuda")
    n = C.c_double()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(reps + 1):
        e0.record()
        _lib.check(L.orp_probe_throughput(kind, iters, _lib.ptr(sink), C.byref(n), _lib.stream_ptr()))
        e1.record()
        torch.cuda.synchronize()
        best = max(best, n.value / (e0.elapsed_time(e1) * 1e-3))
    return best


if __name__ == "__main__":
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    try:
        mhz = float(subprocess.check_output(["nvidia-smi", "--query-gpu=clocks.max.sm", "--format=csv,noheader,nounits"]).split()[0])
    except Exception:
        mhz = 1965.0
    for name, kind, flop in (("FFMA2 (fma.rn.f32x2)", 0, 4 * 32), ("FFMA (3 registers)", 1, 2 * 32), ("FADD", 2, 32), ("LOP3", 3, 0),
                             ("HMMA.1688.F32.TF32", 4, 2 * 16 * 8 * 8)):
        r = probe(kind)
        print(f"{name:24s} {r / (sms * 4 * mhz * 1e6):6.3f} warp instr / cycle / SMSP at {mhz:.0f} MHz"
              + (f"   {r * flop / 1e12:8.2f} TFLOP/s" if flop else ""), flush=True)
