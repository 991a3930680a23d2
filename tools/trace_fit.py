#!/usr/bin/env python
"""Phase breakdown of one optimiser step of the persistent fit kernel (needs the -DORP_FIT_TRACE build:
ORP_B200_LIB=.../liborp_b200_trace.so python tools/trace_fit.py [m] [parts] [peers])."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import orp_b200  # noqa: E402
from orp_b200.fit import ReplicationNet, Optimizer, init_params  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 24
parts = int(sys.argv[2]) if len(sys.argv) > 2 else 8
peers = len(sys.argv) > 3 and sys.argv[3] == "peers"
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 252)
tm = orp_b200.simulate_paths("gbm_log", p, m=m, steps=2, time_major=True)
vals = torch.empty(2 ** m, dtype=torch.float32, device="cuda")
orp_b200.payoff_sums(tm, [100.0], [-1], "call", values_out=vals, value_scale=0.01)
mdl = ReplicationNet("eo", init_params("eo", 1, 0.1), "cuda", peer_exchange=peers)
opt = Optimizer("cuda")
b = mdl.batch(tm[1], tm[2], vals, 0.01, 1.0, 1.0)
for rep in range(2):
    mdl.scratch.zero_()
    mdl.fit(b, opt, epochs=40, patience=1000, use_schedule=False, steps_per_epoch=parts)
torch.cuda.synchronize()
raw = mdl.scratch.cpu().numpy()[-65536:].view(np.uint64)[:2 * 512 * 8].reshape(2, 512, 8).astype(np.float64)
names = ["accumulate", "block total + prefetch", "lr", "own column: poll + sum + publish", "poll totals", "(sync)", "adam + callbacks"]
n_steps = min(512, 40 * parts)
for blk, label in ((0, "block 0"), (1, "block G-1")):
    t = raw[blk, 8:n_steps]                                   # skip the first steps (cold)
    d = np.diff(t, axis=1)
    print(f"{label}: step total {np.mean(t[1:, 0] - t[:-1, 0]) / 1e3:.2f} us  (2^{m} samples / {parts} parts)")
    for k in range(7):
        print(f"   {names[k]:34s} {d[:, k].mean() / 1e3:7.2f} us")
    print(f"   {'step end -> next start':34s} {np.mean(t[1:, 0] - t[:-1, 7]) / 1e3:7.2f} us")
