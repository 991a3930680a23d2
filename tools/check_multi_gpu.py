#!/usr/bin/env python
"""Multi-GPU check (run under torchrun, one rank per GPU): the in-kernel peer exchange of the persistent fit against
the two-stage NCCL form, bit-identical weights on every rank, and the timing of both.

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
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1.0 / D)
tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=D, shard=(rank, world), device=dev, time_major=True)
vals = torch.empty(tm.shape[1], dtype=torch.float32, device=dev)
orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
bond = orp_b200.bond_curve(0.05, 1.0, D + 1)
out = {}
for name, env in (("peers", None), ("nccl", "1")):
    if env is None:
        os.environ.pop("ORP_B200_NO_PEERS", None)
    else:
        os.environ["ORP_B200_NO_PEERS"] = env
    for rep in range(2):
        torch.cuda.synchronize(); dist.barrier()
        t0 = time.perf_counter()
        res = orp_b200.fit_replication(tm, bond, vals.clone(), net="eo", scale=100.0, total_paths=2 ** m)
        torch.cuda.synchronize(); dist.barrier()
        dt = time.perf_counter() - t0
    w = torch.from_numpy(res["params"]).to(dev)
    allw = [torch.empty_like(w) for _ in range(world)]
    dist.all_gather(allw, w)
    same = all(torch.equal(allw[0], x) for x in allw)
    epochs = sum(s["epochs_run"] for s in res["steps"])
    out[name] = (res, dt, same, epochs)
    if rank == 0:
        print(f"{name:6s} world={world} m={m} D={D}: V0={res['V0'] * 100:.6f} phi0={res['phi0']:.6f} epochs={epochs} "
              f"launches={res['launches']} time={dt:.3f}s  weights identical on all ranks: {same}", flush=True)
if rank == 0:
    a, b = out["peers"][0], out["nccl"][0]
    print("peers vs nccl: V0 diff", abs(a["V0"] - b["V0"]), "max |dw|", float(np.abs(a["params"] - b["params"]).max()),
          "epochs", out["peers"][3], out["nccl"][3], flush=True)
    ok = out["peers"][2] and out["nccl"][2] and abs(a["V0"] - b["V0"]) < 2e-2 * abs(b["V0"])
    print("MULTI_GPU_CHECK", "PASS" if ok else "FAIL", flush=True)
dist.destroy_process_group()
