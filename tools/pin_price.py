#!/usr/bin/env python
"""Pins the t = 0 Monte-Carlo price of the full-size configuration against the oracle over the WHOLE path set:
the device payoff sum of 2^m Sobol paths x D steps (BASELINE C5: m = 26, D = 252; discounted mean = the price the
reference prints as 'Discounted Eq[V(T)]', European Options.ipynb raw 3861-3882 / cell 8) against the reference's
NumPy/SciPy path code (oracle.paths, pinned bit-exactly to the reference's arrays) run chunk by chunk on all host cores.
north_star's target is 1e-4 on the price; the sums agree to ~1e-12 and the out-of-the-money counts exactly.

    python tools/pin_price.py [m] [D] [chunk_log2]
"""
import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 26
D = int(sys.argv[2]) if len(sys.argv) > 2 else 252
chunk = 2 ** (int(sys.argv[3]) if len(sys.argv) > 3 else 14)
SIM = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / D)
STRIKES, COLS = [100.0, 90.0, 110.0], [D, D, D // 2]


def _chunk(start):
    os.environ["OMP_NUM_THREADS"] = "1"
    from oracle import paths as opaths
    return opaths.payoff_sums_window("gbm_log", SIM, D + 1, start, min(chunk, 2 ** m - start), STRIKES, COLS, "call")


def main():
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:           # forked before torch / CUDA are touched
        parts = pool.map(_chunk, range(0, 2 ** m, chunk), chunksize=4)
    t_cpu = time.perf_counter() - t0
    sums = np.sum([p[0] for p in parts], axis=0)               # chunk totals added in chunk order (fp64)
    otm = np.sum([p[1] for p in parts], axis=0)
    import torch
    import orp_b200
    t0 = time.perf_counter()
    tm = orp_b200.simulate_paths("gbm_log", SIM, m=m, steps=D, time_major=True)
    dsum, dotm = orp_b200.payoff_sums(tm, STRIKES, COLS, "call")
    dsum, dotm = dsum.cpu().numpy(), dotm.cpu().numpy()
    t_gpu = time.perf_counter() - t0
    disc = np.exp(-0.05 * np.array(COLS) / D)
    out = {"paths_log2": m, "steps": D, "host_cores": cores, "oracle_seconds": t_cpu, "device_seconds_incl_setup": t_gpu,
           "strikes": STRIKES, "columns": COLS,
           "price_oracle": (disc * sums / 2 ** m).tolist(), "price_device": (disc * dsum / 2 ** m).tolist(),
           "rel_diff_sum": (np.abs(dsum - sums) / np.abs(sums)).tolist(), "otm_equal": bool(np.array_equal(dotm, otm)),
           "oracle_path_steps_per_s": 2 ** m * D / t_cpu}
    print(json.dumps(out))
    ok = max(out["rel_diff_sum"]) <= 1e-9 and out["otm_equal"]
    print("PIN_PRICE", "PASS" if ok else "FAIL")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
