"""Drop-in replacement for the reference's `garch_monte_carlo` (app.py:586-657) plus the
north_star's GBM and portfolio modes behind the same return structure.

    from montecarlo_b200 import garch_monte_carlo     # replaces the definition in app.py
    mc_results = garch_monte_carlo(df, mc_sims, forecast_days, current_price, update_status)

Same positional signature (app.py:586-587), same `None` for short histories (app.py:594-595),
same single progress callback (app.py:597-598), same nine result keys (app.py:647-657).
Everything new is keyword-only with reference-equivalent defaults."""
from __future__ import annotations

import functools
import math
import os
from typing import Dict, Optional

import numpy as np

from .spec import (MIN_HISTORY, REF_ALPHA, REF_BETA, REF_OMEGA, REF_PATH_PCTS, REF_TAIL_PCT,
                   REF_TERMINAL_PCTS, PortfolioSpec, SimSpec)

# paths are returned as host NumPy arrays (strict structural parity with app.py:648-649) up to
# this many elements; beyond it they stay on the device as torch tensors
HOST_PATHS_LIMIT = 1 << 26
# full paths are only materialised when (T+1)*N*itemsize fits this budget
DEVICE_PATHS_LIMIT_BYTES = 64 << 30
# when the paths are not kept, the fan chart (app.py:643-645) is computed by regeneration: chunks of paths
# take turns in one device buffer of at most this many bytes
FAN_BUFFER_BYTES = 8 << 30


def _fan_plan(store: bool, want: bool, n_sims: int, n_days: int, dtype: str):
    """-> (store_paths flag for the kernels, fan_chart flag, chunk size for regeneration or 0)."""
    if not want:
        return store, False, 0
    if store:
        return True, True, 0
    itemsize = 8 if dtype == "float64" else 4
    chunk = max(4096, (FAN_BUFFER_BYTES // ((n_days + 1) * itemsize)) // 4 * 4)
    return True, True, int(chunk)


def _column(df, name):
    """df[name] as an array.  pandas builds a Series for `df[name]` (25 us in pandas 3 -- more than the simulation of
    the app's default call takes); the block manager hands out the column's array directly.  Anything unusual
    (no such private accessor, duplicate labels, a non-DataFrame mapping) takes the public route."""
    try:
        return df._get_column_array(df.columns.get_loc(name))
    except Exception:
        return df[name]


def _history_inputs(df):
    """The two DataFrame reads of the reference: app.py:592 and app.py:606."""
    returns = np.asarray(_column(df, "log_ret"), dtype=np.float64)
    nan = np.isnan(returns)
    if nan.any():
        returns = returns[~nan]                        # .dropna()
    rvol_last = float(np.asarray(_column(df, "rvol_20"), dtype=np.float64)[-1])      # .iloc[-1]
    return returns, rvol_last


def _population_std(r: np.ndarray) -> float:
    """np.std(r) for a 1-D float64 array, bit for bit, without its 20 us of dispatch: the very operations of
    numpy/_core/_methods.py:_var (pairwise np.add.reduce, true divide, subtract, multiply in place, reduce, divide,
    sqrt).  tests/test_host_boundary.py holds it to np.std."""
    n = r.shape[0]
    mean = np.add.reduce(r) / n
    x = r - mean
    np.multiply(x, x, out=x)
    return float(np.sqrt(np.add.reduce(x) / n))


def bootstrap_pool(returns: np.ndarray) -> np.ndarray:
    """std_returns (app.py:609): scaled by the population std, not demeaned."""
    if isinstance(returns, np.ndarray) and returns.ndim == 1 and returns.dtype == np.float64 and returns.size:
        return returns / (_population_std(returns) + 1e-10)
    return returns / (np.std(returns) + 1e-10)


def _new_seed() -> int:
    # the reference draws from NumPy's unseeded global generator (no seed control anywhere in app.py)
    return int.from_bytes(os.urandom(8), "little")


PINNED_RESULT_LIMIT_BYTES = 256 << 20


def _to_host(t):
    """Device tensor -> NumPy.  Through page-locked memory from torch's caching host allocator (the array owns the
    block until it is dropped): the copy runs at PCIe speed instead of being staged page by page -- at BASELINE
    configs[0] (20 MB of paths) that is 0.8 ms instead of 5-9."""
    import torch
    if t.numel() * t.element_size() <= PINNED_RESULT_LIMIT_BYTES:
        try:
            host = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            host.copy_(t)
            return host.numpy()
        except RuntimeError:
            pass
    return t.to("cpu").numpy()


@functools.lru_cache(maxsize=16)
def _hist_edges(lo: float, hi: float, bins: int) -> np.ndarray:
    """np.histogram's edges for (lo, hi, bins): the same (read-only) array for repeated calls with the same range."""
    e = np.linspace(lo, hi, bins + 1)
    e.setflags(write=False)
    return e


def _result_dict(out, s0, return_paths: str) -> Dict:
    res = {}
    paths, final = out.paths, out.terminal
    if return_paths == "numpy":
        if paths is not None:
            paths = _to_host(paths).astype(np.float64, copy=False)
            final = paths[-1, :]                       # a view of the last row, like app.py:636
        else:
            final = _to_host(final).astype(np.float64, copy=False)
    res["paths"] = paths
    res["final_prices"] = final
    summ = out.summary()
    res["percentiles"] = {k: np.float64(v) for k, v in summ["percentiles"].items()}
    res["path_percentiles"] = out.fan
    for k in ("expected_return", "expected_vol", "prob_profit"):
        res[k] = np.float64(summ[k])
    for k, v in summ.items():
        if k.startswith("var_") or k.startswith("cvar_"):
            res[k] = np.float64(v)
    if out.hist is not None:
        res["histogram"] = {"counts": out.hist, "range": out.hist_range,
                            "edges": _hist_edges(out.hist_range[0], out.hist_range[1], len(out.hist))}
    res["timing"] = {"sim_ms": out.sim_ms, "total_ms": out.total_ms, "launches": out.launches}
    return res


def _decide_paths(n_sims, n_days, dtype, return_paths):
    itemsize = 8 if dtype == "float64" else 4
    elems = (n_days + 1) * n_sims
    if return_paths in (None, "auto"):
        if elems <= HOST_PATHS_LIMIT:
            return True, "numpy"
        if elems * itemsize <= DEVICE_PATHS_LIMIT_BYTES:
            return True, "device"
        return False, "device"
    if return_paths in (True, "numpy"):
        return True, "numpy"
    if return_paths == "device":
        return True, "device"
    if return_paths in (False, "none"):
        return False, "device"
    raise ValueError("return_paths must be 'auto', 'numpy', 'device' or False")


def garch_monte_carlo(df, n_sims: int, n_days: int, current_price: float, progress_callback=None, *,
                      seed: Optional[int] = None, dtype: str = "float64", return_paths="auto",
                      omega: float = REF_OMEGA, alpha: float = REF_ALPHA, beta: float = REF_BETA,
                      shock: str = "bootstrap", injected_shocks=None,
                      hist_bins: int = 0, hist_range=None, device: Optional[int] = None,
                      group=None, fan_chart: Optional[bool] = None) -> Optional[Dict]:
    """Monte Carlo with volatility clustering: bootstrap shocks + GARCH(1,1) variance,
    exactly the model of app.py:592-657, computed on a B200.

    Extra keyword-only options: `seed` (Philox key; random if None, like the unseeded
    reference), `dtype` ('float64' default = the reference's precision, or 'float32'),
    `return_paths` ('auto' | 'numpy' | 'device' | False), `shock` ('bootstrap' |
    'normal' | 'injected' with `injected_shocks[T, N]`), `hist_bins`/`hist_range`
    (adds a 'histogram' key), `group` (torch.distributed group: shard paths over its ranks),
    `fan_chart` (None: always produce `path_percentiles` -- from the stored paths, or, when the paths
    are not kept, by simulating chunks of them again; False: skip it).
    """
    from .engine import get_engine
    returns, rvol_last = _history_inputs(df)
    if len(returns) < MIN_HISTORY:                     # app.py:594-595
        return None
    if progress_callback:                              # app.py:597-598
        progress_callback("Running Monte Carlo simulation...")
    current_var = rvol_last ** 2 / 252                 # app.py:606
    pool = bootstrap_pool(returns)                     # app.py:609
    return _reference_model(pool, current_var, current_price, n_sims, n_days, seed=seed, dtype=dtype,
                            return_paths=return_paths, omega=omega, alpha=alpha, beta=beta, shock=shock,
                            injected_shocks=injected_shocks, hist_bins=hist_bins, hist_range=hist_range,
                            device=device, group=group, fan_chart=fan_chart)


def _reference_model(pool, current_var, current_price, n_sims, n_days, *, seed, dtype, return_paths, omega, alpha,
                     beta, shock, injected_shocks, hist_bins, hist_range, device, group, fan_chart) -> Dict:
    """app.py:612-657 on the device, from the routine's derived inputs (host array or device tensor pool)."""
    from .engine import get_engine
    n_sims, n_days = int(n_sims), int(n_days)
    store, how = _decide_paths(n_sims, n_days, dtype, return_paths)
    if group is not None and how == "numpy" and return_paths in (None, "auto"):
        how = "device"                                 # shards stay on their home GPU
    # the dashboard reads path_percentiles (app.py:972-981): there by default, whether or not paths are kept
    want_fan = (store or injected_shocks is None) if fan_chart is None else bool(fan_chart)
    keep, fan, chunk = _fan_plan(store, want_fan, n_sims, n_days, dtype)
    spec = SimSpec(n_paths=n_sims, n_steps=n_days, s0=float(current_price), shock=shock, var_model="garch",
                   var0=float(current_var), omega=omega, alpha=alpha, beta=beta, dtype=dtype,
                   seed=_new_seed() if seed is None else int(seed), store_paths=keep, fan_chart=fan,
                   fan_chunk_paths=chunk,
                   pcts=REF_TERMINAL_PCTS, path_pcts=REF_PATH_PCTS, tail_pct=REF_TAIL_PCT,
                   hist_bins=hist_bins, hist_range=hist_range)
    out = get_engine(device).simulate(spec, pool=pool if shock == "bootstrap" else None,
                                      shocks=injected_shocks, group=group, keep_paths=store)
    return _result_dict(out, float(current_price), how)


SMA_WARMUP_ROWS = 199      # close.rolling(200).mean() (app.py:252) is NaN on rows 0..198; app.py:393 drops those rows


def garch_monte_carlo_from_close(close, n_sims: int, n_days: int, progress_callback=None, *, lengths=None,
                                 first_row: int = SMA_WARMUP_ROWS, seed: Optional[int] = None,
                                 dtype: str = "float64", return_paths="auto", omega: float = REF_OMEGA,
                                 alpha: float = REF_ALPHA, beta: float = REF_BETA, hist_bins: int = 0,
                                 hist_range=None, device: Optional[int] = None, fan_chart: Optional[bool] = None,
                                 batched: Optional[bool] = None, calibrate: bool = False, calibrate_levels: int = 8):
    """The routine straight from close prices, without pandas: the device derives `log_ret`,
    `rvol_20`, `std_returns`, `current_var` and `current_price` (app.py:224, 233, 392-393, 592, 606, 609, 1213)
    for every price history in one launch, then runs the model of app.py:612-657 per history.

    `close`: one history (1-D -> one result) or [n_series, n_obs] (-> list of results; `lengths[b]` prices are
    valid in row b).  An entry is None where the reference returns None (fewer than 100 returns left,
    app.py:594-595).  Seeds: history b uses `seed + b`.  `batched` (default: 2-D fp64 input): all histories go
    through ONE simulation launch and one selection per result kind (`mc_run_batch`) instead of one run per
    history; the results are bit-identical either way.  `calibrate=True` (opt-in: it changes the
    results) replaces the reference's hard-coded omega/alpha/beta (app.py:601-603) by each history's own
    quasi-maximum-likelihood values (`mc_calibrate_garch`); every result then carries a 'garch' key.  Raises ValueError for NaN / non-positive prices (the
    reference would silently drop such rows)."""
    from . import _lib as L
    from .engine import get_engine
    single = getattr(close, "ndim", None) == 1 or (hasattr(close, "dim") and close.dim() == 1)
    eng = get_engine(device)
    pool, info = eng.prepare_inputs(close, lengths, first_row, dtype)
    if np.any(info["flags"] & L.INPUT_NAN):
        raise ValueError("close prices must be positive and finite")
    base = _new_seed() if seed is None else int(seed)
    n_sims, n_days = int(n_sims), int(n_days)
    live = [int(n) >= MIN_HISTORY for n in info["pool_len"]]       # app.py:594-595
    if progress_callback:                                          # app.py:597-598, once per history that runs
        for ok in live:
            if ok:
                progress_callback("Running Monte Carlo simulation...")
    fit = None
    if calibrate:
        if dtype != "float64":
            pool64, _ = eng.prepare_inputs(close, lengths, first_row, "float64")
        else:
            pool64 = pool
        fit = eng.calibrate_garch(pool64, info, calibrate_levels, MIN_HISTORY)
    if batched is None:       # the batched run is fp64, no histogram, fan chart from stored paths
        keeps_paths = _decide_paths(n_sims, n_days, dtype, return_paths)[0]
        batched = not single and dtype == "float64" and hist_bins == 0 and (keeps_paths or fan_chart is False)
    if batched:
        if dtype != "float64" or hist_bins:
            raise ValueError("the batched run is fp64 without histogram")
        res = _run_batches(eng, pool, info, n_sims, n_days, base, omega, alpha, beta, return_paths, fan_chart, fit)
        return res[0] if single else res
    results = []
    for b in range(len(info)):
        if not live[b]:
            results.append(None)
            continue
        n = int(info["pool_len"][b])
        o, a, be = (omega, alpha, beta) if fit is None else (float(fit[1][k][b]) for k in ("omega", "alpha", "beta"))
        results.append(_reference_model(pool[b, :n], float(info["var0"][b]), float(info["s0"][b]), n_sims, n_days,
                                        seed=(base + b) & 0xFFFFFFFFFFFFFFFF, dtype=dtype, return_paths=return_paths,
                                        omega=o, alpha=a, beta=be, shock="bootstrap", injected_shocks=None,
                                        hist_bins=hist_bins, hist_range=hist_range, device=device, group=None,
                                        fan_chart=fan_chart))
        if fit is not None:
            results[-1]["garch"] = _garch_entry(fit[1], b)
    return results[0] if single else results


BATCH_PATHS_BYTES = 8 << 30       # histories are simulated together while their stored paths fit this


def _garch_entry(fit, b):
    return {"omega": float(fit["omega"][b]), "alpha": float(fit["alpha"][b]), "beta": float(fit["beta"][b]),
            "objective": float(fit["objective"][b])}


def calibrate_garch(close, *, lengths=None, first_row: int = SMA_WARMUP_ROWS, levels: int = 8,
                    device: Optional[int] = None):
    """GARCH(1,1) parameters of each price history by Gaussian quasi-maximum likelihood under the reference's
    recursion (app.py:632-633), in place of its constants (app.py:601-603).  SURVEY 8f N4.  -> list of
    dict(omega, alpha, beta, objective) / None (fewer than 100 returns); one dict for a 1-D `close`."""
    from .engine import get_engine
    single = getattr(close, "ndim", None) == 1 or (hasattr(close, "dim") and close.dim() == 1)
    eng = get_engine(device)
    pool, info = eng.prepare_inputs(close, lengths, first_row, "float64")
    _, fit = eng.calibrate_garch(pool, info, levels, MIN_HISTORY)
    out = [_garch_entry(fit, b) if fit["valid"][b] else None for b in range(len(fit))]
    return out[0] if single else out


def _batch_dicts(res, fan_all, terminal, paths, info, how, fit):
    """Result dicts (app.py:647-657) for every history of one mc_run_batch, built from the arrays in one go."""
    s0 = info["s0"]
    with np.errstate(all="ignore"):
        expected_return = (res["mean"] / s0 - 1) * 100            # app.py:652
        expected_vol = res["std"] / s0 * 100                      # app.py:653
        prob_profit = res["n_above"] / np.maximum(res["n"], 1) * 100
        tail_k = REF_TERMINAL_PCTS.index(REF_TAIL_PCT)
        var_tail = (s0 - res["pct"][:, tail_k]) / s0 * 100        # app.py:655
        cvar_tail = (s0 - res["tail_sum"] / np.maximum(res["tail_count"], 1)) / s0 * 100   # app.py:656
    names = [f"p{p:g}" for p in REF_TERMINAL_PCTS]
    fan_names = [f"p{p:g}" for p in REF_PATH_PCTS]
    fan_t = np.ascontiguousarray(fan_all.transpose(0, 2, 1)) if fan_all is not None else None
    if how == "numpy":
        host_paths = paths.to("cpu").numpy() if paths is not None else None
        host_final = terminal.to("cpu").numpy() if paths is None else None
    out = []
    for b in range(len(res)):
        if res["n"][b] == 0:
            out.append(None)
            continue
        d = {}
        if how == "numpy":
            d["paths"] = host_paths[b] if host_paths is not None else None
            d["final_prices"] = host_paths[b, -1, :] if host_paths is not None else host_final[b]   # a view, app.py:636
        else:
            d["paths"] = paths[b] if paths is not None else None
            d["final_prices"] = terminal[b]
        d["percentiles"] = dict(zip(names, res["pct"][b, :len(names)]))
        d["path_percentiles"] = dict(zip(fan_names, fan_t[b])) if fan_t is not None else None
        d["expected_return"], d["expected_vol"], d["prob_profit"] = expected_return[b], expected_vol[b], prob_profit[b]
        d[f"var_{100 - REF_TAIL_PCT:g}"], d[f"cvar_{100 - REF_TAIL_PCT:g}"] = var_tail[b], cvar_tail[b]
        d["timing"] = {"sim_ms": float(res["sim_ms"][b]), "total_ms": float(res["total_ms"][b]),
                       "launches": int(res["launches"][b])}
        if fit is not None:
            d["garch"] = _garch_entry(fit, b)
        out.append(d)
    return out


def _run_batches(eng, pool, info, n_sims, n_days, base, omega, alpha, beta, return_paths, fan_chart, fit=None):
    store, how = _decide_paths(n_sims, n_days, "float64", return_paths)
    want_fan = store if fan_chart is None else bool(fan_chart)
    if want_fan and not store:
        raise ValueError("the batched run takes the fan chart from stored paths: pass return_paths='device'")
    per_series = (n_days + 1) * n_sims * 8 if store else n_sims * 8
    step = max(1, min(len(info), BATCH_PATHS_BYTES // per_series))
    results = []
    for b0 in range(0, len(info), step):
        sl = slice(b0, min(len(info), b0 + step))
        res, fan_all, terminal, paths = eng.simulate_batch(
            pool[sl], info[sl], n_sims, n_days, (base + b0) & 0xFFFFFFFFFFFFFFFF, omega=omega, alpha=alpha, beta=beta,
            pcts=REF_TERMINAL_PCTS, path_pcts=REF_PATH_PCTS, tail_pct=REF_TAIL_PCT, store_paths=store,
            fan_chart=want_fan, fit=None if fit is None else (fit[0][sl], fit[1][sl]), raw=True)
        results += _batch_dicts(res, fan_all, terminal, paths, info[sl], how, None if fit is None else fit[1][sl])
    return results


def simulate_gbm(n_sims: int, n_days: int, current_price: float, mu: float, sigma: float, *,
                 dt: float = 1.0 / 252, seed: int = 0, dtype: str = "float32", return_paths=False,
                 fan_chart: Optional[bool] = None, pcts=REF_TERMINAL_PCTS, path_pcts=REF_PATH_PCTS,
                 tail_pct: Optional[float] = REF_TAIL_PCT, hist_bins: int = 0, hist_range="auto",
                 injected_shocks=None, device: Optional[int] = None, group=None, full_resolution: bool = False) -> Dict:
    """Mode G (north_star configs 2/4/5): geometric Brownian motion with normal shocks,
    x += (mu - sigma^2/2) dt + sigma sqrt(dt) z, same reductions and result keys as
    `garch_monte_carlo`.  `hist_range='auto'` = S0*exp(+-6 sigma sqrt(T dt)).

    The float32 shocks are Philox-4x32-10 -> Box-Muller with EIGHT normals per 128-bit call: a 20-bit radius uniform
    and one of 4096 tabulated directions per pair.  Mean and variance are exact to 1e-5, P(|z| > 4) is 0.4 % low,
    P(|z| > 5) 27 % low and |z| never exceeds 5.27 -- far inside the sampling error of any statistic of a multi-step
    path, visible only in the one-step marginal at the 1e-6 level.  `full_resolution=True` draws four normals per
    call from full 32-bit uniforms instead (|z| <= 6.66, directions through MUFU sin/cos; ~25 % slower).
    float64 always draws from 32-bit uniforms."""
    from .engine import get_engine
    store, how = _decide_paths(n_sims, n_days, dtype, return_paths)
    if hist_bins and hist_range == "auto":
        w = 6.0 * sigma * math.sqrt(n_days * dt)
        hist_range = (current_price * math.exp(-w), current_price * math.exp(w))
    elif hist_range == "auto":
        hist_range = None
    want_fan = store if fan_chart is None else bool(fan_chart)      # GBM default: only from stored paths
    keep, fan, chunk = _fan_plan(store, want_fan, int(n_sims), int(n_days), dtype)
    spec = SimSpec.gbm(int(n_sims), int(n_days), float(current_price), mu, sigma, dt, dtype=dtype, seed=int(seed),
                       shock="injected" if injected_shocks is not None else "normal", store_paths=keep,
                       fan_chart=fan, fan_chunk_paths=chunk,
                       pcts=tuple(pcts), path_pcts=tuple(path_pcts), tail_pct=tail_pct,
                       hist_bins=hist_bins, hist_range=hist_range, full_res_normals=bool(full_resolution))
    out = get_engine(device).simulate(spec, shocks=injected_shocks, group=group, keep_paths=store)
    return _result_dict(out, float(current_price), how)


def simulate_portfolio(n_sims: int, n_days: int, s0, mu, sigma, weights, corr=None, *, chol=None,
                       dt: float = 1.0 / 252, seed: int = 0, dtype: str = "float32", return_paths=False,
                       pcts=REF_TERMINAL_PCTS, path_pcts=REF_PATH_PCTS, tail_pct: Optional[float] = REF_TAIL_PCT,
                       hist_bins: int = 0, hist_range=None, injected_shocks=None,
                       device: Optional[int] = None, group=None) -> Dict:
    """Mode P (north_star config 3): d correlated assets, y = L z with L = chol(corr);
    statistics are those of the portfolio value V = sum_i w_i S_i (VaR/CVaR of V)."""
    from .engine import get_engine
    store, how = _decide_paths(n_sims, n_days, dtype, return_paths)
    pf = PortfolioSpec(s0=s0, mu=mu, sigma=sigma, weights=weights, corr=corr, chol=chol, dt=dt)
    v0 = pf.arrays()["v0"]
    spec = SimSpec(n_paths=int(n_sims), n_steps=int(n_days), s0=v0,
                   shock="injected" if injected_shocks is not None else "normal", var_model="const",
                   dtype=dtype, seed=int(seed), store_paths=store, fan_chart=store, pcts=tuple(pcts),
                   path_pcts=tuple(path_pcts), tail_pct=tail_pct, hist_bins=hist_bins, hist_range=hist_range,
                   portfolio=pf)
    out = get_engine(device).simulate(spec, shocks=injected_shocks, group=group)
    return _result_dict(out, v0, how)
