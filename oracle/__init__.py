"""CPU oracle for the Monte Carlo hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and there only as the checker or the timed CPU arm.
The shipped path (``montecarlo_b200``) never imports this package and raises if
its CUDA library is missing.

What the oracle restates (reference = /root/reference/app.py):

* ``mc_numpy``   -- the NumPy loop of ``garch_monte_carlo`` (app.py:586-657) and
                    the north_star's GBM / Cholesky-portfolio extensions, which
                    have no counterpart in the reference.
* ``philox``     -- bit model of the on-device generator (Philox-4x32-10, the
                    uniform bit-stuffing, Box-Muller, bootstrap index).
* ``select_model`` -- NumPy stage backend (exact radix select) for the multi-rank stage protocol.
* ``ref_loader`` -- AST extraction of the unmodified reference function; usable
                    only where /root/reference exists (this container), used by
                    ``make_golden.py`` to pin the restatement.

Parity status: Mode R (bootstrap + GARCH) is PINNED by fixtures generated from
the unmodified reference (tests/golden/ref_*.npz, oracle/make_golden.py).
Modes G and P, and the histogram, do not exist in the reference: "parity
unpinned" by the reference; they are pinned by closed-form lognormal answers and
Philox known-answer vectors instead.
"""
