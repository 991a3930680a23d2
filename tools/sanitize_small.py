#!/usr/bin/env python
"""A few small launches of every kernel, for compute-sanitizer (memcheck / racecheck, one tool per run)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "option-replicating-portfolio-with-neural-networks_b200")]
import torch  # noqa: E402
import orp_b200  # noqa: E402
from orp_b200.fit import ReplicationNet, Optimizer, init_params  # noqa: E402

dev = torch.device("cuda:0")
p = dict(S0=100.0, mu=0.05, sigma=0.2, dt=1 / 12, s0=0.16679, a=0.0033, b=0.156, c=0.0158)
x = orp_b200.sobol_uint32(10, 7, 1235, device=dev)
for model, order, dtype, keep in (("gbm_log", "incremental", "float64", 1), ("gbm_log", "incremental", "float32", 3),
                                  ("gbm_arith", "incremental", "float64", 1), ("sv", "incremental", "float64", 1),
                                  ("gbm_log", "bridge", "float64", 1)):
    tm = orp_b200.simulate_paths(model, p, start=3, count=4099, steps=12, ordering=order, dtype=dtype, keep_every=keep,
                                 device=dev, time_major=True)
tm = orp_b200.simulate_paths("gbm_log", p, m=13, steps=12, device=dev, time_major=True)
vals = torch.empty(2 ** 13, dtype=torch.float32, device=dev)
orp_b200.payoff_sums(tm, [90.0, 100.0, 110.0], [12, 6, 12], "call", values_out=vals, value_scale=0.01)
bond = orp_b200.bond_curve(0.05, 1.0, 13)
for net in ("eo", "rp"):
    for persistent, two_stage, peers in ((True, False, False), (False, False, False), (False, True, False), (True, False, True)):
        mdl = ReplicationNet(net, init_params(net, 1, 0.1 if net == "eo" else [0.9, 0.1]), dev, peer_exchange=peers)
        mdl.persistent, mdl.two_stage = persistent, two_stage
        opt = Optimizer(dev)
        b = mdl.batch(tm[11], tm[12], vals, 0.01, 1.0, 1.0)
        mdl.fit(b, opt, epochs=3, patience=2, use_schedule=True, steps_per_epoch=3)
        mdl.predict(b, torch.empty_like(vals))
        mdl.close()
mdl = ReplicationNet("eo", init_params("eo", 1, 0.1), dev)
mdl.fit(mdl.batch(tm[11], tm[12], vals, 0.01, 1.0, 1.0), Optimizer(dev), epochs=1, patience=2, use_schedule=True, mode="keras")
torch.cuda.synchronize()
print("sanitize_small ok")
