#!/usr/bin/env python
"""Benchmark of the B200 hot path on BASELINE.json's scaling configuration (configs[4], "C5"):

    2^26 scrambled-Sobol paths x 252 steps, log-Euler GBM (S0 = K = 100, T = 1, r = .05, sigma = .2), fp64,
    every column materialised, terminal call payoff, full backward replication (252 fits of the 97-parameter
    net, scaled mode), sharded over N GPUs by contiguous Sobol index ranges (strong scaling: 2^26 paths in total).

One "step" = one pass of the whole pipeline (paths -> payoff -> 252 fits -> V0, phi0, psi0).
`value` = paths x steps / step time (path-steps/s), whole job, max over ranks.  The JSON line also carries
the per-stage times, the HBM roofline of the path kernel, the fp32 figure of the fit kernel, the end-to-end
number through the public API with host-side set-up and H2D/D2H inside the timed region, the sampled clocks
and a CPU baseline (the reference's NumPy/SciPy path code + the torch-CPU restatement of its Keras fit).

    python bench.py [--gpus N] [--steps K] [--warmup W]          # own arm
    python bench.py --impl reference ...                         # the reference's CPU path, all host threads
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "path-steps/sec"
CFG = dict(S0=100.0, K=100.0, T=1.0, r=0.05, sigma=0.2, D=252)
FP32_NOMINAL_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12     # SIMT fp32 FMA peak at max clock (MEASURED_PEAKS.json has no fp32 figure:
                                                          # the run measures its own with orp_probe_throughput, see fp32_peak_measured)
FIT_FLOP_PER_SAMPLE = 2 * 232 + 56                        # 232 multiply-adds + 56 elementwise ops per sample (EO net, fwd+bwd)
FIT_FMA_PIPE_SLOTS = 144 * 2                              # FMA-pipe issue cycles per warp-sample: 144 packed FFMA2/FMUL2/FADD2-class ops, 2 cycles each
FIT_PEAK_SAMPLES_PER_S = 148 * 4 * 32 * 1.965e9 / FIT_FMA_PIPE_SLOTS     # what the FMA pipe allows at max clock


def fp32_peak_measured():
    """Packed-FMA (fma.rn.f32x2) peak of THIS GPU in TFLOP/s, measured in this process with the library's probe kernel
    (csrc/orp_probe.cu; best of 5 launches of ~10 ms, CUDA events): the denominator of the fit kernel's roofline."""
    import ctypes as C
    import torch
    from orp_b200 import _lib
    L = _lib.lib()
    sink = torch.zeros(4, dtype=torch.float32, device="cuda")
    n = C.c_double()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(6):
        e0.record()
        _lib.check(L.orp_probe_throughput(0, 2000, _lib.ptr(sink), C.byref(n), _lib.stream_ptr()))
        e1.record()
        torch.cuda.synchronize()
        best = max(best, n.value / (e0.elapsed_time(e1) * 1e-3))
    return best * 32 * 4 / 1e12                 # warp instructions/s x 32 lanes x 2 FMAs x 2 flop


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0}, "fallback"


def profiled_traffic(m, D):
    """dram bytes (read + write) of one paths_kernel launch of THIS workload from the newest committed `ncu --set full`
    summary (profiles/r0N_paths_kernel_traffic.json), else None."""
    for name in ("r02_paths_kernel_traffic.json", "r01_paths_kernel_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                t = json.load(f)
            if t.get("paths_log2") == m and t.get("path_steps") == D:
                return t["dram_bytes_read"] + t["dram_bytes_write"]
        except (OSError, KeyError, ValueError):
            pass
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) > 3 + k and r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ own arm

class Pipeline:
    """One rank's share of the C5 pipeline; buffers are allocated once and reused by every step."""

    def __init__(self, m, D, rank, world, device):
        import torch
        import orp_b200
        self.torch, self.orp = torch, orp_b200
        self.m, self.D, self.rank, self.world, self.device = m, D, rank, world, device
        self.start, self.count = orp_b200.shard_range(2 ** m, rank, world)
        self.sim = dict(S0=CFG["S0"], mu=CFG["r"], sigma=CFG["sigma"], dt=CFG["T"] / D)
        self.paths = torch.empty((D + 1, self.count), dtype=torch.float64, device=device)
        self.vals = torch.empty(self.count, dtype=torch.float32, device=device)
        self.bond = orp_b200.bond_curve(CFG["r"], CFG["T"], D + 1)
        self.ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        # C4-style sweep on the same path set: 8 strikes x 8 maturity columns = 64 European calls in one payoff call
        self.sweep_k = np.repeat(np.linspace(80.0, 120.0, 8), 8)
        self.sweep_c = np.tile(np.minimum(np.arange(1, 9) * (D // 8) + D % 8, D), 8)

    def sweep(self, reps=5):
        """Options priced per second (north_star's second throughput figure): 64 (strike, maturity) pairs, (a) off this
        rank's STORED paths (payoff_partial + payoff_final kernels) and (b) straight from the Sobol engine with the fused
        generate + payoff kernel that never stores a path (C4 as BASELINE words it: generation included).  Device time,
        one all-reduce of the packed sums per call included."""
        torch, orp = self.torch, self.orp
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        kw = dict(r=CFG["r"], total_paths=2 ** self.m)
        orp.price_european(self.paths, self.sweep_k, self.sweep_c, "call", dt=CFG["T"] / self.D, **kw)
        e[0].record()
        for _ in range(reps):
            prices, _ = orp.price_european(self.paths, self.sweep_k, self.sweep_c, "call", dt=CFG["T"] / self.D, **kw)
        e[1].record()
        fkw = dict(steps=self.D, start=self.start, count=self.count, device=self.device, **kw)
        orp.price_european_fused("gbm_log", self.sim, self.sweep_k, self.sweep_c, "call", **fkw)
        e[2].record()
        for _ in range(reps):
            fused, _ = orp.price_european_fused("gbm_log", self.sim, self.sweep_k, self.sweep_c, "call", **fkw)
        e[3].record()
        torch.cuda.synchronize()
        return 64 * reps / (e[0].elapsed_time(e[1]) * 1e-3), prices, 64 * reps / (e[2].elapsed_time(e[3]) * 1e-3), fused

    def step(self, fresh_engine=False):
        torch, orp = self.torch, self.orp
        if fresh_engine:                       # e2e: host-side set-up and the H2D of its tables are inside the timed region
            from orp_b200 import sobol
            sobol.engine.cache_clear()
        self.ev[0].record()
        orp.simulate_paths("gbm_log", self.sim, steps=self.D, start=self.start, count=self.count, out=self.paths,
                           device=self.device, time_major=True)
        self.ev[1].record()
        sums, otm = orp.payoff_sums(self.paths, [CFG["K"]], [-1], "call", values_out=self.vals, value_scale=1.0 / CFG["S0"])
        self.ev[2].record()
        res = orp.fit_replication(self.paths, self.bond, self.vals, net="eo", scale=CFG["S0"], total_paths=2 ** self.m)
        self.ev[3].record()
        est = torch.stack([sums[0], otm[0]])
        orp.dist.allreduce_sum_(est)
        est = est.cpu().numpy()                # D2H of the step's result
        self.ev[3].synchronize()
        t = [self.ev[i].elapsed_time(self.ev[i + 1]) for i in range(3)]
        return dict(price_mc=float(np.exp(-CFG["r"] * CFG["T"]) * est[0] / 2 ** self.m), p_otm=float(est[1] / 2 ** self.m),
                    V0=res["V0"] * CFG["S0"], phi0=res["phi0"], psi0=res["psi0"], launches=3 + res["launches"],
                    epochs=sum(s["epochs_run"] for s in res["steps"]), t_paths_ms=t[0], t_payoff_ms=t[1], t_fit_ms=t[2])


def cpu_baseline(m_sample, D, threads=1, batch_size=512, start=0):
    """The reference's CPU path on a bounded sample of the same workload: NumPy/SciPy path code (oracle.paths,
    a line-for-line restatement pinned bit-exactly to the reference's arrays) + payoff + the torch-CPU
    restatement of the Keras loop (batch 512, as the reference runs it)."""
    import torch
    from oracle import fit as ofit, paths as opaths
    torch.set_num_threads(threads)
    n = 2 ** m_sample
    t0 = time.perf_counter()
    sim = dict(S0=CFG["S0"], mu=CFG["r"], sigma=CFG["sigma"], dt=CFG["T"] / D)
    if threads > 1 and n >= 4 * 4096:
        import multiprocessing as mp
        chunks = [(s, min(4096, n - s)) for s in range(0, n, 4096)]
        with mp.get_context("fork").Pool(min(threads, len(chunks))) as pool:
            S = np.concatenate(pool.starmap(_cpu_chunk, [(sim, D, s, c) for s, c in chunks]), axis=0)
    else:
        S = opaths.simulate_window("gbm_log", sim, D + 1, start, n)
    t_paths = time.perf_counter() - t0
    pay = opaths.payoff(S[:, -1], CFG["K"], "call")
    bond = opaths.bond_curve(CFG["r"], CFG["T"], D + 1)
    res = ofit.backward_replication(S, bond, pay / CFG["S0"], net="eo", scale=CFG["S0"], batch_size=batch_size)
    t = time.perf_counter() - t0
    epochs = int(sum(h["epochs_run"] for h in res["history"]))
    return dict(value=n * D / t, seconds=t, seconds_paths=t_paths, V0=res["V0"] * CFG["S0"], epochs=epochs,
                adam_steps=epochs * -(-n // batch_size),
                price_mc=float(np.exp(-CFG["r"] * CFG["T"]) * pay.mean()),
                sample=f"2^{m_sample} of the 2^26 paths x {D} steps, same pipeline (paths + payoff + {D} fits, batch {batch_size})")


def _cpu_chunk(sim, D, start, count):
    from oracle import paths as opaths
    return opaths.simulate_window("gbm_log", sim, D + 1, start, count)


def _cpu_replica(m_sample, D, k):
    r = cpu_baseline(m_sample, D, threads=1, start=k * 2 ** m_sample)
    return r["seconds"], r["V0"], r["price_mc"], r["epochs"], r["adam_steps"]


def run_reference(args):
    """The reference's CPU code on ALL host cores.  Its fit is a serial chain of 512-sample Adam steps on 8-wide layers:
    intra-op threads only add overhead (8 torch threads: 27.7k path-steps/s against 104k with one), so the only way it
    can use more than one core is one process per core, each running the whole pipeline (paths -> payoff -> 252 fits) on
    its own window of 2^ref_sample_m consecutive Sobol paths — every step of this arm is `cores` such bounded samples
    running side by side, and `value` is their aggregate throughput.  The one-core figure is reported next to it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    cores = max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    D = args.path_steps
    n = 2 ** args.ref_sample_m
    times, last = [], None
    with mp.get_context("fork").Pool(cores) as pool:          # forked before this process has run any torch op
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            last = pool.starmap(_cpu_replica, [(args.ref_sample_m, D, k) for k in range(cores)])
            if i >= args.warmup:
                times.append(time.perf_counter() - t0)
    ms = 1e3 * float(np.mean(times))
    value = cores * n * D / (ms / 1e3)
    one_core = n * D / float(np.mean([r[0] for r in last]))   # per-replica rate with every core busy
    sample = (f"{cores} windows of 2^{args.ref_sample_m} of the 2^{args.m} paths x {D} steps side by side, one process per core, "
              f"each the same pipeline (paths + payoff + {D} fits, batch 512)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "path-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": config_dict(args.m, D, args.gpus),
            "cpu_baseline": {"value": value, "unit": "path-steps/s", "cores": cores, "kind": "port", "sample": sample,
                             "host_cores": os.cpu_count(), "per_core_path_steps_per_s": one_core,
                             "note": "the reference's fit is serial (512-sample Adam steps on 8-wide layers; intra-op threads slow "
                                     "it down), so all cores are used the only way it allows: independent processes, one window "
                                     "of paths each"},
            "e2e": {"value": value, "unit": "path-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "check": {"V0": float(np.mean([r[1] for r in last])), "price_mc": float(np.mean([r[2] for r in last])),
                      "epochs_per_window": float(np.mean([r[3] for r in last])), "adam_steps_per_window": float(np.mean([r[4] for r in last])),
                      "samples_per_adam_step": 512, "windows": cores, "paths_per_window": n,
                      "note": "work per path: the reference's Keras schedule (batch 512, shuffled) runs epochs_per_window epochs of "
                              "paths_per_window / 512 Adam steps on each window; the GPU arm runs its epochs (fit_kernel.epochs_per_step) "
                              "of 8 Adam steps over all 2^m paths"}}
    print(json.dumps(line))
    return 0


def sweep_err(prices, pipe):
    from math import erf, exp, log, sqrt
    def bs(K, T):
        N = lambda x: 0.5 * (1.0 + erf(x / sqrt(2.0)))
        d1 = (log(CFG["S0"] / K) + (CFG["r"] + 0.5 * CFG["sigma"] ** 2) * T) / (CFG["sigma"] * sqrt(T))
        return CFG["S0"] * N(d1) - K * exp(-CFG["r"] * T) * N(d1 - CFG["sigma"] * sqrt(T))
    ref = np.array([bs(k, c * CFG["T"] / pipe.D) for k, c in zip(pipe.sweep_k, pipe.sweep_c)])
    return float(np.abs(np.asarray(prices) - ref).max())


def config_dict(m, D, world):
    """The `config` object of BOTH arms (the reference arm times a bounded sample of this workload, see its cpu_baseline)."""
    return {"workload": workload_name(m, D), "paths_log2": m, "path_steps": D,
            "parallelism": f"paths sharded x{world} by Sobol index range",
            "l2": "inputs larger than L2 (the path array is read back by the fits; >= 16 GB per GPU)"}


def workload_name(m, D):
    return (f"C5: 2^{m} Sobol paths x {D} steps log-Euler GBM fp64 (S0=K=100,T=1,r=.05,sigma=.2), all columns stored, "
            f"call payoff, full backward replication ({D} fits, EO net, scaled mode)")


def run_own(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    m, D = args.m, args.path_steps
    pipe = Pipeline(m, D, rank, world, device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        last = pipe.step()
    sampler = ClockSampler(local)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    barrier()
    if rank == 0:
        sampler.start()
    ev[0].record()
    stages = []
    for _ in range(args.steps):
        last = pipe.step()
        stages.append(last)
    ev[1].record()
    barrier()
    elapsed = ev[0].elapsed_time(ev[1]) * 1e-3                 # device time of exactly `steps` steps (CUDA events)
    clocks = sampler.stop() if rank == 0 else None
    # end to end through the public API: host-side engine set-up (NumPy bit stream + integer scramble) + H2D of its
    # tables + pipeline + D2H of the results, every step
    barrier()
    n_e2e = min(args.steps, 10)                               # every timed step up to ten (a C5 step takes ~12 s on one GPU)
    ev[2].record()
    for _ in range(n_e2e):
        pipe.step(fresh_engine=True)
    ev[3].record()
    barrier()
    e2e_elapsed = ev[2].elapsed_time(ev[3]) * 1e-3 / n_e2e
    opts_per_s, sweep_prices, fused_per_s, fused_prices = pipe.sweep()
    fp32_peak = fp32_peak_measured()
    tt = torch.tensor([elapsed, e2e_elapsed, 1.0 / opts_per_s, 1.0 / fused_per_s, -fp32_peak], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    elapsed, e2e_elapsed, inv_opts, inv_fused, neg_peak = (float(x) for x in tt.cpu())
    fp32_peak = -neg_peak                                      # the slowest GPU's packed-FMA peak
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    ms = 1e3 * elapsed / args.steps
    total_steps = 2 ** m * D
    peaks, peak_kind = measured_peaks()
    t_paths = float(np.mean([s["t_paths_ms"] for s in stages]))
    t_fit = float(np.mean([s["t_fit_ms"] for s in stages]))
    t_pay = float(np.mean([s["t_payoff_ms"] for s in stages]))
    path_bytes = pipe.count * (D + 1) * 8                      # algorithmic bytes: one fp64 store per kept column (incl. column 0)
    achieved = path_bytes / (t_paths * 1e-3) / 1e9
    epochs = last["epochs"]
    fit_flops = 2 ** m / world * epochs * FIT_FLOP_PER_SAMPLE
    from orp_b200 import sobol
    eng = sobol.engine(D + 1, 1235)
    line = {
        "metric": METRIC, "value": total_steps / (ms / 1e3), "unit": "path-steps/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_dict(m, D, world),
        "roofline": {"kernel": "paths_kernel<GBM_LOG,double>", "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"],
                     "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": profiled_traffic(m, D) if world == 1 else None,
                     "peak_kind": peak_kind + " (burst copy figure)",
                     "algorithmic_bytes_per_launch": path_bytes, "ms_per_launch": t_paths,
                     "note": "north_star's HBM-bound kernel; by TIME the step is dominated by the fit kernel, which is bound by the "
                             "fp32 FMA pipe (width-8 layers: SIMT by north_star's own rule), see fit_kernel"},
        "fit_kernel": {"kernel": "fit_run_kernel<EO,MSE,float> (persistent, one launch per Model.fit; fp32 price columns in network units)",
                       "bound": "fp32 FMA pipe (packed FFMA2), not hbm/tensor: width-8 layers",
                       "tflops": fit_flops / (t_fit * 1e-3) / 1e12, "peak_measured_tflops": fp32_peak,
                       "frac_of_measured_fp32_peak": fit_flops / (t_fit * 1e-3) / 1e12 / fp32_peak,
                       "peak_how": "fma.rn.f32x2 register loop on every SM, measured in this run (orp_probe_throughput)",
                       "peak_nominal_tflops": FP32_NOMINAL_TFLOPS,
                       "epochs_per_step": epochs, "samples_per_s_per_gpu": 2 ** m / world * epochs / (t_fit * 1e-3),
                       "fma_pipe_peak_samples_per_s": FIT_PEAK_SAMPLES_PER_S,
                       "frac_of_fma_pipe": 2 ** m / world * epochs / (t_fit * 1e-3) / FIT_PEAK_SAMPLES_PER_S,
                       "note": "whole fit stage incl. predict passes and launch gaps"},
        "stages_ms": {"paths": t_paths, "payoff": t_pay, "fit": t_fit},
        "path_stage_path_steps_per_s": total_steps / world / (t_paths * 1e-3) * world,
        "e2e": {"value": total_steps / e2e_elapsed, "unit": "path-steps/s", "h2d_bytes_per_step": int(eng.sv32.nbytes + eng.tlane.nbytes),
                "d2h_bytes_per_step": int(16 + 16 * 8 * (D + 2) + 32 * D)},
        "options_priced_per_s": {"value": 1.0 / inv_opts, "what": "64 European calls (8 strikes x 8 maturities) off the stored path set of this "
                                 "workload, payoff_partial + payoff_final kernels + all-reduce (C4-shaped sweep)",
                                 "max_abs_err_vs_black_scholes": sweep_err(sweep_prices, pipe),
                                 "fused_generate_payoff": {"value": 1.0 / inv_fused, "what": "the same 64 calls straight from the Sobol engine: "
                                                           "paths_payoff_kernel generates the 2^%d x %d path-steps, pays off and reduces without storing a path "
                                                           "(0 B of path traffic)" % (m, D),
                                                           "equivalent_path_steps_per_s": total_steps / (64 * inv_fused),
                                                           "equivalent_GBps": 8 * total_steps / (64 * inv_fused) / 1e9,
                                                           "max_rel_diff_vs_stored_sweep": float(np.abs(np.asarray(fused_prices) / np.asarray(sweep_prices) - 1).max())}},
        "gpu_launches": int(sum(s["launches"] for s in stages)),
        "clocks": clocks,
        "result": {"V0_nn": last["V0"], "phi0": last["phi0"], "psi0": last["psi0"], "price_mc": last["price_mc"], "p_otm": last["p_otm"]},
    }
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(args.cpu_sample_m, D, threads=1)
        line["cpu_baseline"] = {"value": cb["value"], "unit": "path-steps/s", "cores": 1, "kind": "port", "sample": cb["sample"],
                                "seconds": cb["seconds"], "paths_only_path_steps_per_s": 2 ** args.cpu_sample_m * D / cb["seconds_paths"]}
        from oracle import paths as opaths
        line["result"]["black_scholes"] = opaths.black_scholes_call(CFG["S0"], CFG["K"], CFG["T"], CFG["r"], CFG["sigma"])
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--m", type=int, default=26, help="log2 of the total path count (BASELINE C5: 26)")
    ap.add_argument("--path-steps", type=int, default=CFG["D"], help="time steps per path (BASELINE C5: 252)")
    ap.add_argument("--cpu-sample-m", type=int, default=12)
    ap.add_argument("--ref-sample-m", type=int, default=12)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    return run_reference(args) if args.impl == "reference" else run_own(args)


if __name__ == "__main__":
    sys.exit(main())
