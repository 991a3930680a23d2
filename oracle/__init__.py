"""CPU oracle for the B200 hot path — TEST INFRASTRUCTURE, not product code.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this package.  The product
(``option-replicating-portfolio-with-neural-networks_b200/``) never imports it
and has no CPU fallback.

What it restates (NumPy / SciPy / torch-CPU), with the reference lines each
module follows:

* ``sobol``   – scrambled Sobol set-up and points: ``scipy.stats.qmc.Sobol`` as
                called at ``Replicating_Portfolio.py:54-57`` (SciPy 1.18.1,
                ``scipy/stats/_qmc.py:1761-1828, 1862-1917``).
* ``paths``   – GBM (arithmetic ``Replicating_Portfolio.py:64-65`` and log-Euler
                ``European Options.ipynb`` cell 6), the stochastic-volatility
                recursion ``Replicating_Portfolio.py:280-289``, bond curve
                ``:68-69``, column reduction ``:92-96``, payoffs ``:88-89`` and
                ``European Options.ipynb`` cell 8.
* ``bridge``  – Brownian-bridge ordering (NEW feature, no reference counterpart:
                parity unpinned by the reference, anchored on analytic checks).
* ``fit``     – torch-CPU restatement of the Keras fit loop
                ``Replicating_Portfolio.py:128-227`` / ``European Options.ipynb``
                cells 12-13 (TensorFlow is not installed here, so fit parity is
                pinned only statistically — see DESIGN.md).

Parity pinning: ``sobol`` is checked bit-for-bit against SciPy's own ``_sv`` /
``_shift`` / points; ``paths`` is checked against fixtures produced by running the
UNMODIFIED reference sources from ``/root/reference`` (``tests/golden/make_golden.py``)
and against the numbers printed in the reference's saved notebook outputs.
"""
