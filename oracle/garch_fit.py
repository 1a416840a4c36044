"""GARCH(1,1) calibration by Gaussian quasi-maximum likelihood -- CPU restatement (TEST INFRASTRUCTURE).

SURVEY 8f N4.  The reference has NO calibration: omega, alpha, beta are hard-coded locals (app.py:601-603) of the
recursion var' = omega + alpha r^2 + beta var (app.py:632-633).  Parity is therefore "unpinned by the
reference"; this module restates the algorithm of montecarlo_b200/csrc/mc_calibrate.cu operation by operation
(same objective, same variance targeting, same 16 x 16 grid refinement), and `tests/test_calibrate.py` anchors it
on what CAN be checked independently: a general-purpose optimiser (scipy Nelder-Mead) on the same objective never
finds a better likelihood, and the parameters of simulated GARCH data are recovered.
"""
from __future__ import annotations

import numpy as np

GRID = 16
ALPHA_MAX = 0.5
PERSIST_MIN = 0.5
PERSIST_MAX = 0.9995


def objective(z2, alpha, beta, s2=None):
    """sum_t [log h_t + z_t^2 / h_t], h_0 = s2 = mean(z^2), omega = s2 (1 - alpha - beta).  alpha/beta may be
    arrays (candidates evaluated side by side)."""
    z2 = np.asarray(z2, dtype=np.float64)
    s2 = float(np.mean(z2)) if s2 is None else s2
    alpha, beta = np.broadcast_arrays(np.asarray(alpha, dtype=np.float64), np.asarray(beta, dtype=np.float64))
    omega = s2 * (1.0 - (alpha + beta))
    h = np.full(alpha.shape, s2)
    total = np.zeros(alpha.shape)
    for x in z2:
        total += np.log(h) + x / h
        h = omega + alpha * x + beta * h
    return total


def calibrate(returns, levels: int = 8):
    """-> dict(omega, alpha, beta, objective) for one history of raw log returns (app.py:592)."""
    r = np.asarray(returns, dtype=np.float64)
    c = np.std(r) + 1e-10                        # the pool's scale, app.py:609
    z2 = (r / c) ** 2
    s2 = float(np.mean(z2))
    a_lo, a_step = 0.0, ALPHA_MAX / (GRID - 1)
    p_lo, p_step = PERSIST_MIN, (PERSIST_MAX - PERSIST_MIN) / (GRID - 1)
    ia, ip = np.meshgrid(np.arange(GRID), np.arange(GRID))          # candidate index = ip * GRID + ia
    best = None
    for _ in range(levels):
        alpha = a_lo + ia * a_step
        persist = p_lo + ip * p_step
        beta = persist - alpha
        ok = (alpha >= 0.0) & (beta >= 0.0) & (persist <= PERSIST_MAX + 1e-12) & (persist > 0.0)
        val = np.full(alpha.shape, np.inf)
        with np.errstate(all="ignore"):
            # evaluated exactly as the kernel does: omega from the persistence, not from alpha + beta
            omega = s2 * (1.0 - persist[ok])
            h = np.full(omega.shape, s2)
            tot = np.zeros(omega.shape)
            for x in z2:
                tot += np.log(h) + x / h
                h = omega + alpha[ok] * x + beta[ok] * h
        val[ok] = tot
        val = np.where(np.isnan(val), np.inf, val)
        k = int(np.argmin(val.ravel()))                             # ties: lowest candidate index
        wa, wp = k % GRID, k // GRID
        best = (a_lo + wa * a_step, p_lo + wp * p_step, float(val.ravel()[k]))
        a_lo, a_step = best[0] - 2.0 * a_step, 4.0 * a_step / (GRID - 1)
        p_lo, p_step = best[1] - 2.0 * p_step, 4.0 * p_step / (GRID - 1)
    alpha, persist, value = best
    return dict(alpha=alpha, beta=persist - alpha, omega=s2 * (1.0 - persist) * c * c, objective=value, s2=s2, scale=c)


def simulate_garch(n, omega, alpha, beta, seed=0):
    """Returns of a GARCH(1,1) process with Gaussian innovations (for the recovery test)."""
    rs = np.random.RandomState(seed)
    h = omega / (1.0 - alpha - beta)
    out = np.empty(n)
    for t in range(n):
        out[t] = np.sqrt(h) * rs.standard_normal()
        h = omega + alpha * out[t] ** 2 + beta * h
    return out
