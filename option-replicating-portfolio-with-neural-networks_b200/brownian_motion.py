"""Drop-in for the reference's ``brownian_motion.py`` (same names, signatures and results).

``get_dW`` / ``get_W`` draw a single length-``T`` vector from NumPy's legacy global generator; that is a
sequential host stream, not data-parallel work, so these two stay on the host and reproduce the
reference bit for bit (``brownian_motion.py:6-24``; pinned by ``tests/golden/brownian.npz``).

The Brownian-BRIDGE ordering BASELINE.json's north_star attributes to this file does not exist in the
reference.  It is provided here as an additive API on top of the GPU path (:func:`bridge_paths`).
"""
from __future__ import annotations

from typing import Optional

import numpy as np


def get_dW(T: int, random_state: Optional[int] = None) -> np.ndarray:
    """``T`` standard-normal increments of a Brownian motion; ``random_state`` seeds NumPy's global
    legacy generator exactly as the reference does."""
    np.random.seed(random_state)
    return np.random.normal(0.0, 1.0, T)


def get_W(T: int, random_state: Optional[int] = None) -> np.ndarray:
    """Brownian motion sampled at unit steps, starting at 0: ``[0, dW_0, dW_0+dW_1, ...]`` (length ``T``)."""
    steps = get_dW(T, random_state)
    path = np.empty(T, dtype=steps.dtype)
    path[:1] = 0.0
    np.cumsum(steps[:-1], out=path[1:])
    return path


# ---- additive API (not in the reference) -----------------------------------------------------------

def bridge_paths(m: int, steps: int, params: dict, model: str = "gbm_log", seed: int = 1235, **kw):
    """Sobol paths with Brownian-bridge time ordering on the GPU; see ``orp_b200.simulate_paths``."""
    from orp_b200 import simulate_paths
    return simulate_paths(model, params, m=m, steps=steps, seed=seed, ordering="bridge", **kw)


def bridge_schedule(steps: int):
    """The bisection program (Sobol dimension, weights) used by the bridge kernels."""
    from orp_b200 import bridge_schedule_host
    return bridge_schedule_host(steps)
