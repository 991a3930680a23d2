#!/usr/bin/env python
"""Multi-GPU check (run under torchrun, one rank per GPU).  Asserts, and prints one PASS / FAIL line per item:

  1. the in-kernel peer exchange of the persistent fit against the two-stage form with a collective between the
     stages: bit-identical weights, and the same weights on every rank;
  2. RANK-COUNT INVARIANCE on real hardware: rank 0 alone (a process group of one) runs the same 2^m paths on its
     GPU; V0, phi0 and every fitted weight must equal the sharded run's bit for bit;
  3. the sharded VaR quantiles (radix select with the ranks' histograms summed) equal the one-GPU quantiles;
  4. the sharded device survivors equal the unsharded array (counter-based draws indexed by the global path).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/check_multi_gpu.py [m] [D]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import orp_b200  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 22
D = int(sys.argv[2]) if len(sys.argv) > 2 else 12
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
solo = dist.new_group([0])                       # rank 0 on its own: the one-GPU reference run
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / D)
bond = orp_b200.bond_curve(0.05, 1.0, D + 1)
ok_all = True


def report(name, ok, detail=""):
    global ok_all
    ok_all = ok_all and bool(ok)
    if rank == 0:
        print(f"{'PASS' if ok else 'FAIL'}  {name}  {detail}", flush=True)


def paths_and_values(shard, device):
    tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, shard=shard, device=device, time_major=True)
    vals = torch.empty(tm.shape[1], dtype=torch.float32, device=device)
    orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
    return tm, vals


tm, vals = paths_and_values((rank, world), dev)
out = {}
for name, env in (("peers", None), ("two-stage", "1")):
    if env is None:
        os.environ.pop("ORP_B200_NO_PEERS", None)
    else:
        os.environ["ORP_B200_NO_PEERS"] = env
    for rep in range(2):
        torch.cuda.synchronize(); dist.barrier()
        t0 = time.perf_counter()
        res = orp_b200.fit_replication(tm, bond, vals.clone(), net="eo", scale=100.0, total_paths=2 ** m, var_quantiles=(.98, .99))
        torch.cuda.synchronize(); dist.barrier()
        dt = time.perf_counter() - t0
    w = torch.from_numpy(res["params"]).to(dev)
    allw = [torch.empty_like(w) for _ in range(world)]
    dist.all_gather(allw, w)
    same = all(torch.equal(allw[0], x) for x in allw)
    epochs = sum(s["epochs_run"] for s in res["steps"])
    out[name] = (res, dt, same, epochs)
    if rank == 0:
        print(f"{name:9s} world={world} m={m} D={D}: V0={res['V0'] * 100:.12f} phi0={res['phi0']:.12f} epochs={epochs} "
              f"launches={res['launches']} time={dt:.3f}s", flush=True)
    report(f"{name}: weights identical on all {world} ranks", same)
a, b = out["peers"][0], out["two-stage"][0]
report("peer exchange == two-stage form, bit for bit",
       np.array_equal(a["params"], b["params"]) and a["V0"] == b["V0"] and out["peers"][3] == out["two-stage"][3],
       f"max |dw| = {float(np.abs(a['params'] - b['params']).max()):.3g}")

# 2./3.: the same job on ONE GPU (rank 0, group of one)
if rank == 0:
    tm1, vals1 = paths_and_values((0, 1), dev)
    one = orp_b200.fit_replication(tm1, bond, vals1, net="eo", scale=100.0, total_paths=2 ** m, group=solo, var_quantiles=(.98, .99))
    print(f"one GPU   m={m} D={D}: V0={one['V0'] * 100:.12f} phi0={one['phi0']:.12f} "
          f"epochs={sum(s['epochs_run'] for s in one['steps'])}", flush=True)
    report(f"rank-count invariance: {world} GPUs == 1 GPU, bit for bit (V0, phi0, psi0, weights, epochs per fit)",
           one["V0"] == a["V0"] and one["phi0"] == a["phi0"] and one["psi0"] == a["psi0"] and np.array_equal(one["params"], a["params"])
           and [s["epochs_run"] for s in one["steps"]] == [s["epochs_run"] for s in a["steps"]],
           f"V0 {one['V0']!r} vs {a['V0']!r}")
    report("sharded VaR quantiles == one-GPU quantiles at every step",
           all(s1["VaR"] == sN["VaR"] for s1, sN in zip(one["steps"], a["steps"])))
    del tm1, vals1
dist.barrier()

# 4.: survivors
lam_p = dict(S0=0.01, mu=0.075, sigma=0.000597, dt=1.0 / D)
start, count = orp_b200.shard_range(2 ** 16, rank, world)
lam = orp_b200.simulate_paths("linear", lam_p, steps=D, start=start, count=count, seed=1234, device=dev, time_major=True)
mine = orp_b200.survivors(lam, 10_000, 1.0 / D, start=start)
parts = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(parts, mine)
if rank == 0:
    lam1 = orp_b200.simulate_paths("linear", lam_p, steps=D, start=0, count=2 ** 16, seed=1234, device=dev, time_major=True)
    whole = orp_b200.survivors(lam1, 10_000, 1.0 / D, start=0)
    report("sharded survivors == unsharded array", torch.equal(torch.cat(parts, dim=1), whole))
flag = torch.tensor([1.0 if ok_all else 0.0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("MULTI_GPU_CHECK", "PASS" if float(flag) == 1.0 else "FAIL", flush=True)
dist.destroy_process_group()
