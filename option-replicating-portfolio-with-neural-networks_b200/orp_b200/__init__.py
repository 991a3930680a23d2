"""orp_b200 — B200 (sm_100a) implementation of the QMC replicating-portfolio hot path of
ithakis/Option-Replicating-Portfolio-with-Neural-Networks.

Python host layer over the C ABI of ``liborp_b200.so`` (``include/orp_b200.h``).  The reference-compatible
entry points live one directory up as ``Replicating_Portfolio.py`` / ``brownian_motion.py``; this package
holds the additive API the BASELINE configurations need.  No CPU fallback anywhere.
"""
from .sobol import SobolEngine, sobol_uint32, sobol_normal, ndtri_u32          # noqa: F401
from .paths import (simulate_paths, payoff_sums, price_european, bond_curve,   # noqa: F401
                    n_time_steps_of, reduction_of, shard_range, bridge_schedule_host,
                    payoff_sums_fused, price_european_fused)
from .fit import fit_replication, ReplicationNet, Optimizer, init_params, n_params, scale_column, release_peers   # noqa: F401
from .risk import quantiles, survivors                                                    # noqa: F401
from .calibrate import CIRParams, estimate_CIR_params                          # noqa: F401
from . import dist                                                             # noqa: F401

__all__ = ["SobolEngine", "sobol_uint32", "sobol_normal", "ndtri_u32", "simulate_paths", "payoff_sums",
           "price_european", "payoff_sums_fused", "price_european_fused", "bond_curve", "n_time_steps_of", "reduction_of", "shard_range",
           "bridge_schedule_host", "quantiles", "survivors", "CIRParams", "estimate_CIR_params", "fit_replication", "ReplicationNet", "Optimizer", "init_params", "n_params",
           "dist"]
