"""NumPy restatement of the reference's Monte Carlo routine (TEST INFRASTRUCTURE).

Reference: /root/reference/app.py:586-657 (`garch_monte_carlo`).  Each function
cites the lines it follows.  Modes G (normal shocks, constant sigma) and P
(Cholesky-correlated portfolio) are north_star extensions with NO counterpart in
the reference ("parity unpinned" by the reference; pinned by closed forms).

State layout mirrors the reference: time-major ``paths[(T+1), N]`` (app.py:612)
and one variance per path per step (app.py:615), float64 throughout.
"""
from __future__ import annotations

import numpy as np

REF_OMEGA = 0.00001   # app.py:601
REF_ALPHA = 0.10      # app.py:602
REF_BETA = 0.85       # app.py:603
REF_TERMINAL_PCTS = (1, 5, 10, 25, 50, 75, 90, 95, 99)   # app.py:639
REF_PATH_PCTS = (5, 25, 50, 75, 95)                       # app.py:644
MIN_HISTORY = 100                                         # app.py:594


def bootstrap_pool(returns):
    """std_returns = returns / (std + 1e-10): scaled, NOT demeaned, ddof=0 (app.py:609)."""
    r = np.asarray(returns, dtype=np.float64)
    return r / (r.std() + 1e-10)


SMA_WARMUP_ROWS = 199     # close.rolling(200).mean() (app.py:252) is NaN on rows 0..198; app.py:393 drops them


def history_inputs(close, first_row=SMA_WARMUP_ROWS):
    """What the routine reads, from a close-price history, with the reference's own pandas
    calls: log_ret (app.py:224), rvol_20 (app.py:233), rows >= first_row kept (app.py:392-393), then
    returns (app.py:592), current_var (app.py:606), std_returns (app.py:609), current_price (app.py:1213).
    Returns None when nothing is left, else dict(returns, pool, pool_std, var0, s0, rvol_last)."""
    import pandas as pd
    close = pd.Series(np.asarray(close, dtype=np.float64))
    log_ret = np.log(close / close.shift(1))
    rvol_20 = log_ret.rolling(20).std() * np.sqrt(252)
    frame = pd.DataFrame({"Close": close, "log_ret": log_ret, "rvol_20": rvol_20})
    frame = frame.replace([np.inf, -np.inf], 0).iloc[first_row:].dropna()
    if len(frame) == 0:
        return None
    returns = frame["log_ret"].dropna().values
    rvol_last = float(frame["rvol_20"].iloc[-1])
    std = np.std(returns)
    return dict(returns=returns, pool=returns / (std + 1e-10), pool_std=float(std), rvol_last=rvol_last,
                var0=initial_variance(rvol_last), s0=float(frame["Close"].iloc[-1]))


def initial_variance(rvol_last):
    """annualised vol -> daily variance (app.py:606)."""
    return float(rvol_last) ** 2 / 252


def step_paths(shocks, s0, var0, omega=REF_OMEGA, alpha=REF_ALPHA, beta=REF_BETA,
               garch=True, drift=0.0):
    """Evolve all paths through the given shock matrix ``shocks[T, N]``.

    garch=True : app.py:618-633 -- vol = sqrt(var); r = shock*vol; S *= exp(r);
                 var' = omega + alpha r^2 + beta var.  `drift` (per step, log
                 space) is an extension; the reference has none.
    garch=False: constant variance var0 (Mode G): r = drift + shock*sqrt(var0).
    Returns float64 paths[(T+1), N].
    """
    z = np.asarray(shocks, dtype=np.float64)
    n_steps, n_paths = z.shape
    paths = np.empty((n_steps + 1, n_paths))
    paths[0] = s0
    var = np.full(n_paths, float(var0))
    for t in range(n_steps):
        innov = z[t] * np.sqrt(var)               # app.py:623-626
        paths[t + 1] = paths[t] * np.exp(innov + drift)   # app.py:629
        if garch:
            var = omega + alpha * (innov * innov) + beta * var   # app.py:632-633
    return paths


def terminal_statistics(final, s0, pcts=REF_TERMINAL_PCTS):
    """The scalar reductions of app.py:636-640 and app.py:652-656 on float64 values."""
    x = np.asarray(final, dtype=np.float64)
    pct = {f"p{p}": np.percentile(x, p) for p in pcts}
    out = {
        "percentiles": pct,
        "expected_return": (x.mean() / s0 - 1) * 100,
        "expected_vol": x.std() / s0 * 100,
        "prob_profit": (x > s0).mean() * 100,
    }
    if "p5" in pct:
        tail = x[x <= pct["p5"]]
        out["var_95"] = (s0 - pct["p5"]) / s0 * 100
        out["cvar_95"] = (s0 - tail.mean()) / s0 * 100
    return out


def fan_chart(paths, pcts=REF_PATH_PCTS):
    """Per-time-step percentiles across paths (app.py:643-645)."""
    p = np.asarray(paths, dtype=np.float64)
    return {f"p{q}": np.percentile(p, q, axis=1) for q in pcts}


def summarize(paths, s0, pcts=REF_TERMINAL_PCTS, path_pcts=REF_PATH_PCTS):
    """Return dict with the reference's nine keys (app.py:647-657)."""
    final = paths[-1]
    out = {"paths": paths, "final_prices": final, "path_percentiles": fan_chart(paths, path_pcts)}
    out.update(terminal_statistics(final, s0, pcts))
    return out


def garch_monte_carlo_port(returns, rvol_last, n_sims, n_days, s0, *, shocks=None,
                           indices=None, rng=None):
    """Port of the whole routine on plain arrays (the DataFrame reads of
    app.py:592/606 are done by the caller).  Shock source, in priority order:
    `shocks[T,N]` values, `indices[T,N]` into the pool, or `rng` (a
    numpy RandomState; draws like app.py:620)."""
    returns = np.asarray(returns, dtype=np.float64)
    if len(returns) < MIN_HISTORY:              # app.py:594-595
        return None
    pool = bootstrap_pool(returns)
    if shocks is None:
        if indices is not None:
            shocks = pool[np.asarray(indices)]
        else:
            rng = rng if rng is not None else np.random
            shocks = np.stack([rng.choice(pool, size=n_sims) for _ in range(n_days)])
    paths = step_paths(shocks, s0, initial_variance(rvol_last))
    return summarize(paths, s0)


# --------------------------------------------------------------------------- Mode G
def gbm_step_params(mu, sigma, dt):
    """(drift per step, variance per step) of the log-Euler GBM step."""
    return (mu - 0.5 * sigma * sigma) * dt, sigma * sigma * dt


def gbm_paths(z, s0, mu, sigma, dt):
    drift, var = gbm_step_params(mu, sigma, dt)
    return step_paths(z, s0, var, garch=False, drift=drift)


def gbm_closed_form(s0, mu, sigma, horizon, q=None):
    """Lognormal known answers: mean, std, and quantiles of S_T."""
    from scipy.stats import norm
    mean = s0 * np.exp(mu * horizon)
    std = mean * np.sqrt(np.expm1(sigma * sigma * horizon))
    out = {"mean": mean, "std": std}
    if q is not None:
        qq = np.asarray(q, dtype=np.float64)
        out["quantiles"] = s0 * np.exp((mu - 0.5 * sigma ** 2) * horizon
                                       + sigma * np.sqrt(horizon) * norm.ppf(qq))
    return out


# --------------------------------------------------------------------------- Mode P
def portfolio_terminal(z, s0, mu, sigma, chol, weights, dt):
    """z[T, d, N] iid normals; y = L z; per-asset log-Euler step; V = sum w_i S_i(T).
    Returns (V[N], S_T[d, N])."""
    z = np.asarray(z, dtype=np.float64)
    n_steps, d, n_paths = z.shape
    s0 = np.broadcast_to(np.asarray(s0, dtype=np.float64), (d,))
    mu = np.broadcast_to(np.asarray(mu, dtype=np.float64), (d,))
    sigma = np.broadcast_to(np.asarray(sigma, dtype=np.float64), (d,))
    L = np.asarray(chol, dtype=np.float64)
    x = np.zeros((d, n_paths))
    drift = (mu - 0.5 * sigma ** 2) * dt
    scale = sigma * np.sqrt(dt)
    for t in range(n_steps):
        y = L @ z[t]
        x += drift[:, None] + scale[:, None] * y
    s_t = s0[:, None] * np.exp(x)
    v = np.asarray(weights, dtype=np.float64) @ s_t
    return v, s_t


# --------------------------------------------------------------------------- histogram
def histogram(values, bins, lo, hi):
    """np.histogram on float64-cast values: the semantics row H of SURVEY 8a pins."""
    counts, edges = np.histogram(np.asarray(values, dtype=np.float64), bins=bins, range=(lo, hi))
    return counts.astype(np.int64), edges
