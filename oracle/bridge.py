"""Oracle: Brownian-bridge time ordering (TEST INFRASTRUCTURE).

NEW feature asked for by BASELINE.json's north_star; the reference has no bridge anywhere
(``brownian_motion.py`` is a plain ``cumsum`` of pseudo-random normals), so this oracle is
**parity unpinned** by the reference and is anchored on analytic properties instead
(``Cov(W_s, W_t) = min(s, t)``, prices converge to Black-Scholes) in ``tests/test_oracle_paths.py``.

Specification (the CUDA path must agree with it):

* The engine is unchanged (dimension ``d = D+1``, same seed): column ``k`` of the normal matrix,
  ``k = 1..D``, is consumed by the ``k``-th bridge node in CONSTRUCTION ORDER instead of by time
  step ``k``; column 0 stays unused.
* Construction order = breadth-first bisection: node ``D`` first (``W_D = sqrt(D) z_1``), then the
  queue of intervals ``(l, r)`` starting with ``(0, D)``: an interval with ``r - l >= 2`` creates
  node ``m = (l + r) // 2`` and appends ``(l, m)``, ``(m, r)``.
* ``W_m = ((r-m) W_l + (m-l) W_r) / (r-l) + sqrt((m-l)(r-m)/(r-l)) z_k`` in units of ``sqrt(dt)``
  (``W_0 = 0``).
* The step loops are unchanged and read ``Z_t = W_t - W_{t-1}`` where they read ``W1[:, t]``.
"""
from __future__ import annotations

from collections import deque

import numpy as np


def bridge_schedule(D: int):
    """Returns int arrays ``(node, left, right, dim)`` of length ``D`` in construction order;
    ``dim[k-1] = k``.  ``left/right`` are the already-built neighbours (``left = 0`` means W=0;
    for the first node ``left = right = 0`` by convention)."""
    node, left, right = [D], [0], [0]
    q = deque([(0, D)])
    while q:
        l, r = q.popleft()
        if r - l >= 2:
            m = (l + r) // 2
            node.append(m); left.append(l); right.append(r)
            q.append((l, m)); q.append((m, r))
    assert len(node) == D and sorted(node) == list(range(1, D + 1))
    return (np.array(node), np.array(left), np.array(right), np.arange(1, D + 1))


def bridge_weights(D: int):
    """Per construction step ``(wl, wr, sd)`` as float64."""
    node, left, right, _ = bridge_schedule(D)
    wl = np.zeros(D); wr = np.zeros(D); sd = np.zeros(D)
    sd[0] = np.sqrt(float(D))
    for k in range(1, D):
        m, l, r = int(node[k]), int(left[k]), int(right[k])
        wl[k] = (r - m) / (r - l)
        wr[k] = (m - l) / (r - l)
        sd[k] = np.sqrt((m - l) * (r - m) / (r - l))
    return wl, wr, sd


def bridge_normals(W: np.ndarray) -> np.ndarray:
    """Map the ``(paths, D+1)`` normal matrix to bridge-ordered unit increments of the same
    shape (column 0 copied through, unused)."""
    D = W.shape[1] - 1
    node, left, right, dim = bridge_schedule(D)
    wl, wr, sd = bridge_weights(D)
    Wt = np.zeros((W.shape[0], D + 1))
    with np.errstate(invalid="ignore"):
        Wt[:, D] = sd[0] * W[:, 1]
        for k in range(1, D):
            m, l, r = int(node[k]), int(left[k]), int(right[k])
            Wt[:, m] = wl[k] * Wt[:, l] + wr[k] * Wt[:, r] + sd[k] * W[:, dim[k]]
        Z = np.empty_like(W)
        Z[:, 0] = W[:, 0]
        Z[:, 1:] = Wt[:, 1:] - Wt[:, :-1]
    return Z
