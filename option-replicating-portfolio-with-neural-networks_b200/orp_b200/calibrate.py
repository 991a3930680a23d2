"""Calibration of the volatility process from a volatility series: the step BEFORE the hot path.

Host-side NumPy, like the reference (``Extra: Stochastic Volatility.ipynb`` cells 3-4): an ordinary-least-squares fit of
``d sigma / sqrt(sigma) = ab / sqrt(sigma) - a sqrt(sigma) + c eps``.  The reference downloads ^GSPC with yfinance and
uses scikit-learn's ``LinearRegression(fit_intercept=False)``; here the series is an argument (no network on the box) and
the regression is ``numpy.linalg.lstsq`` — the same normal equations.  Its outputs ``(a, b, c)`` and the last volatility
``s0`` are the inputs of ``simulate_paths('sv', ...)`` / ``Replicating_Portfolio_SV``.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class CIRParams:
    """CIR process parameters; ``2ab >= c^2`` is required (the reference's check and message, cell 3)."""

    a: float  # mean reversion parameter
    b: float  # asymptotic mean
    c: float  # Brownian motion scale factor (standard deviation)

    def __post_init__(self):
        if 2 * self.a * self.b < self.c ** 2:
            raise ValueError("2ab has to be less than or equal to c^2.")


def estimate_CIR_params(sigma_t: np.ndarray) -> CIRParams:
    """OLS estimate of the CIR parameters from a 1-D volatility series (cell 4 of the Extra notebook)."""
    sigma_t = np.asarray(sigma_t, dtype=np.float64)
    if sigma_t.ndim != 1 or sigma_t.size < 3:
        raise ValueError("sigma_t must be a 1-D series with at least three observations")
    sigma_sqrt = np.sqrt(sigma_t[:-1])
    y = np.diff(sigma_t) / sigma_sqrt
    X = np.stack([1.0 / sigma_sqrt, sigma_sqrt], axis=1)
    coef, *_ = np.linalg.lstsq(X, y, rcond=None)
    a = -coef[1]
    b = coef[0] / a
    c = np.std(y - X @ coef)
    return CIRParams(float(a), float(b), float(c))
