"""GPU parity, part 3: the drop-in function itself, rank-sharding invariance (ranks emulated
on one GPU through the staged C entry points), and size-independent properties at
BASELINE.json's full sizes."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from oracle import mc_numpy as port
from oracle import philox
from montecarlo_b200 import SimSpec, garch_monte_carlo, simulate_gbm, simulate_portfolio
from montecarlo_b200 import _lib as L
from montecarlo_b200.spec import make_params, shard_range

pytestmark = pytest.mark.gpu


def _frame(n_rows=1250, seed=0):
    import pandas as pd
    rs = np.random.RandomState(seed)
    return pd.DataFrame({"log_ret": rs.normal(0.0, 0.02, n_rows), "rvol_20": np.full(n_rows, 0.02 * np.sqrt(252))})


def test_drop_in_contract():
    """Every expression the dashboard evaluates on mc_results (app.py:944-1000)."""
    df, msgs = _frame(), []
    res = garch_monte_carlo(df, 1000, 20, 100.0, msgs.append, seed=1)
    assert msgs == ["Running Monte Carlo simulation..."]                       # app.py:597-598
    assert set(res) >= {"paths", "final_prices", "percentiles", "path_percentiles", "expected_return",
                        "expected_vol", "prob_profit", "var_95", "cvar_95"}     # app.py:647-657
    assert isinstance(res["paths"], np.ndarray) and res["paths"].dtype == np.float64
    assert res["paths"].shape == (21, 1000) and res["final_prices"].shape == (1000,)
    assert np.shares_memory(res["final_prices"], res["paths"])                  # a view, like app.py:636
    assert np.all(res["paths"][0] == 100.0)
    assert list(res["percentiles"]) == [f"p{p}" for p in (1, 5, 10, 25, 50, 75, 90, 95, 99)]
    for k in ("p5", "p25", "p50", "p75", "p95"):
        assert len(res["path_percentiles"][k]) == 21                            # app.py:977
    for k in ("expected_return", "prob_profit", "var_95", "cvar_95"):
        f"{res[k]:+.2f}"                                                        # app.py:950-953
    # against NumPy on the returned arrays: the reference's own reductions
    exp = port.summarize(res["paths"], 100.0)
    for k in ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95"):
        assert abs(res[k] - exp[k]) < 1e-9 * max(1.0, abs(exp[k])), k
    for k, v in exp["percentiles"].items():
        assert res["percentiles"][k] == v
    for k, v in exp["path_percentiles"].items():
        assert np.array_equal(res["path_percentiles"][k], v)
    # reproducible under a seed, different without
    again = garch_monte_carlo(df, 1000, 20, 100.0, seed=1)
    assert np.array_equal(again["paths"], res["paths"])
    assert not np.array_equal(garch_monte_carlo(df, 1000, 20, 100.0)["paths"], res["paths"])
    assert garch_monte_carlo(_frame(99), 1000, 20, 100.0) is None               # app.py:594-595


def test_statistical_parity_with_reference_c1(golden_dir):
    """BASELINE config 1 (10k x 252): device RNG vs the reference run under np.random.seed(42).  north_star (b):
    mean, std and percentiles within 3 standard errors (of the difference of two independent samples) -- the twelve
    scalar comparisons below.  The fan chart adds 5 x 252 strongly correlated comparisons: at 3 SE each, three to four
    would exceed the bound by chance alone, so its bound is on the maximum (4.5 SE) and on the rms (2 SE)."""
    import os
    g = np.load(os.path.join(golden_dir, "ref_c1_seed42.npz"))
    import pandas as pd
    df = pd.DataFrame({"log_ret": g["returns"], "rvol_20": float(g["rvol_last"])})
    n, s0 = int(g["n_sims"]), float(g["s0"])
    res = garch_monte_carlo(df, n, int(g["n_days"]), s0, seed=7, dtype="float32", return_paths="numpy")
    vol = float(g["expected_vol"])                       # std of terminal / s0 * 100
    se_mean = vol / math.sqrt(n) * math.sqrt(2)
    assert abs(res["expected_return"] - float(g["expected_return"])) < 3 * se_mean
    assert abs(res["expected_vol"] / vol - 1) < 3 * math.sqrt(2) * 1.5 / math.sqrt(n)      # heavy-ish tails
    assert abs(res["prob_profit"] - float(g["prob_profit"])) < 3 * 100 * math.sqrt(0.5 / n)
    ref_p = g["percentiles"]
    final = np.sort(res["final_prices"])
    for i, p in enumerate(port.REF_TERMINAL_PCTS):
        # SE of a quantile via the empirical density around it
        q = p / 100
        k = int(q * (n - 1))
        w = max(50, int(0.01 * n))
        dens = (2 * w / n) / (final[min(n - 1, k + w)] - final[max(0, k - w)])
        se = math.sqrt(q * (1 - q) / n) / dens * math.sqrt(2)
        assert abs(res["percentiles"][f"p{p}"] - ref_p[i]) < 3 * se, p
    # fan chart: every time step's five quantiles within 4 SE of the difference of two samples
    fan_ref = g["path_percentiles"]
    rows = np.sort(res["paths"][1:], axis=1)                 # [T, n]
    w = max(50, int(0.01 * n))
    for i, p in enumerate(port.REF_PATH_PCTS):
        q = p / 100
        k = int(q * (n - 1))
        dens = (2 * w / n) / (rows[:, min(n - 1, k + w)] - rows[:, max(0, k - w)])
        se = math.sqrt(q * (1 - q) / n) / dens * math.sqrt(2)
        dev = np.abs(res["path_percentiles"][f"p{p}"][1:] - fan_ref[i][1:]) / se
        assert dev.max() < 4.5, (p, float(dev.max()))         # 1,260 (correlated) comparisons
        assert np.sqrt(np.mean(dev ** 2)) < 2.0, p


def _staged_emulated_ranks(spec, pool, world):
    """Run `world` rank-shards on ONE GPU, one handle per rank, gathering stage buffers by hand
    (kernels are plain sequential launches; nothing waits on another rank)."""
    from montecarlo_b200.engine import Engine, _CudaStageBackend
    engines = [Engine(0) for _ in range(world)]
    dt = torch.float64 if spec.dtype == "float64" else torch.float32
    backs, outs = [], []
    for r, e in enumerate(engines):
        off, cnt = shard_range(spec.n_paths, r, world)
        pool_t = e.to_device(pool, dt) if pool is not None else None
        if pool_t is not None:
            spec.pool_len = pool_t.numel()
        params = make_params(spec, path_offset=off, n_local=cnt, n_ranks=world)
        term = torch.empty(cnt, dtype=dt, device=e.device)
        backs.append(_CudaStageBackend(e, params, None, pool_t, None, None, term, world))
        outs.append(term)
    for b in backs:
        b.begin()
    done = False
    while not done:
        words = backs[0].stage_words()
        assert all(b.stage_words() == words for b in backs)
        gathered = torch.cat([b.stage_view(words) for b in backs])
        done = all([b.next(gathered, world) for b in backs])
    results = []
    for e in engines:
        res = L.McResult()
        hist = np.zeros(spec.hist_bins, dtype=np.uint64)
        e.check(e.lib.mc_fetch_result(e.handle, C.byref(res), C.c_void_p(hist.ctypes.data), None, e.stream_ptr()))
        results.append((res, hist))
    for e in engines:
        e.close()
    return results, torch.cat(outs)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("mode", ["bootstrap", "gbm"])
def test_shard_invariance_emulated_ranks(engine, world, mode):
    n, t = 50_001, 24
    pool = port.bootstrap_pool(np.random.RandomState(4).normal(0, 0.02, 1250)) if mode == "bootstrap" else None
    def mk():
        if mode == "bootstrap":
            return SimSpec(n_paths=n, n_steps=t, s0=100.0, var0=4e-4, seed=11, hist_bins=128, hist_range=(50.0, 200.0))
        return SimSpec.gbm(n, t, 100.0, 0.05, 0.2, seed=11, hist_bins=128, hist_range=(50.0, 200.0))
    single = engine.simulate(mk(), pool=pool)
    results, terminal = _staged_emulated_ranks(mk(), pool, world)
    assert torch.equal(terminal, single.terminal)                 # per-path values independent of the sharding
    for res, hist in results:                                    # every rank ends with the same merged result
        assert res.n == n and res.n_above == single.n_above
        assert [res.pct[i] for i in range(9)] == list(single.pct.values())
        assert res.tail_count == single.tail_count
        assert abs(res.tail_sum - single.tail_sum) <= 1e-12 * abs(single.tail_sum)
        assert abs(res.mean - single.mean) <= 1e-13 * abs(single.mean)
        assert abs(res.std - single.std) <= 1e-11 * single.std
        assert res.min == single.min and res.max == single.max
        assert np.array_equal(hist.astype(np.int64), single.hist)


def test_full_size_c2_properties():
    """BASELINE config 2: 10M x 252 fp32 GBM.  Size-independent checks: exact percentiles of the
    returned terminals, histogram mass, closed-form lognormal answers within 4 SE."""
    n, t, s0, mu, sigma, dt = 10_000_000, 252, 100.0, 0.05, 0.2, 1 / 252
    res = simulate_gbm(n, t, s0, mu, sigma, dt=dt, seed=0x123456789ABCDEF0, hist_bins=2048)
    term = res["final_prices"].cpu().numpy().astype(np.float64)
    assert term.shape == (n,)
    for p, v in zip(port.REF_TERMINAL_PCTS, np.percentile(term, port.REF_TERMINAL_PCTS)):
        assert res["percentiles"][f"p{p}"] == v
    h = res["histogram"]
    assert np.array_equal(h["counts"], np.histogram(term, bins=2048, range=h["range"])[0])
    cf = port.gbm_closed_form(s0, mu, sigma, t * dt, q=np.array(port.REF_TERMINAL_PCTS) / 100)
    assert abs(term.mean() - cf["mean"]) < 4 * cf["std"] / math.sqrt(n)
    assert abs(term.std() / cf["std"] - 1) < 4 * 2.5 / math.sqrt(n)
    from scipy.stats import norm
    for p, qv in zip(port.REF_TERMINAL_PCTS, cf["quantiles"]):
        q = p / 100
        dens = norm.pdf(norm.ppf(q)) / (qv * sigma * math.sqrt(t * dt))
        assert abs(res["percentiles"][f"p{p}"] - qv) < 4 * math.sqrt(q * (1 - q) / n) / dens, p
    exp = port.terminal_statistics(term, s0)
    for k in ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95"):
        assert abs(res[k] - exp[k]) < 1e-9 * max(1.0, abs(exp[k])), k


def test_full_size_c4_share_properties():
    """BASELINE config 4: one GPU's share of 100M x 2520 (12.5M paths, fp32 GBM, ten years of daily steps):
    exact percentiles / histogram / tail on the returned terminals, closed-form lognormal answers within 4 SE."""
    n, t, s0, mu, sigma, dt = 12_500_000, 2520, 100.0, 0.05, 0.2, 1 / 252
    res = simulate_gbm(n, t, s0, mu, sigma, dt=dt, seed=0x123456789ABCDEF0, hist_bins=2048)
    term = res["final_prices"].cpu().numpy().astype(np.float64)
    assert term.shape == (n,)
    for p, v in zip(port.REF_TERMINAL_PCTS, np.percentile(term, port.REF_TERMINAL_PCTS)):
        assert res["percentiles"][f"p{p}"] == v
    h = res["histogram"]
    assert np.array_equal(h["counts"], np.histogram(term, bins=2048, range=h["range"])[0])
    cf = port.gbm_closed_form(s0, mu, sigma, t * dt, q=np.array(port.REF_TERMINAL_PCTS) / 100)
    assert abs(term.mean() - cf["mean"]) < 4 * cf["std"] / math.sqrt(n)
    # std of a lognormal with sigma sqrt(T) = 0.63: the fourth moment is large; SE from the sample itself
    m2 = (term - term.mean()) ** 2
    assert abs(term.std() ** 2 - cf["std"] ** 2) < 4 * m2.std() / math.sqrt(n)
    from scipy.stats import norm
    for p, qv in zip(port.REF_TERMINAL_PCTS, cf["quantiles"]):
        q = p / 100
        dens = norm.pdf(norm.ppf(q)) / (qv * sigma * math.sqrt(t * dt))
        assert abs(res["percentiles"][f"p{p}"] - qv) < 4 * math.sqrt(q * (1 - q) / n) / dens, p
    exp = port.terminal_statistics(term, s0)
    for k in ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95"):
        assert abs(res[k] - exp[k]) < 1e-9 * max(1.0, abs(exp[k])), k


def test_full_size_c3_portfolio_properties():
    """BASELINE config 3: 16 correlated assets, 1M x 252: portfolio mean and VaR/CVaR consistency."""
    d, n, t, dt = 16, 1_000_000, 252, 1 / 252
    corr = np.full((d, d), 0.3) + 0.7 * np.eye(d)
    sigma = 0.15 + 0.01 * np.arange(d)
    res = simulate_portfolio(n, t, np.full(d, 100.0), 0.05, sigma, np.full(d, 1 / d), corr, dt=dt, seed=3)
    v = res["final_prices"].cpu().numpy().astype(np.float64)
    assert abs(v.mean() - 100.0 * math.exp(0.05)) < 4 * v.std() / math.sqrt(n)
    # variance of the log of a single-asset-equivalent is not closed form; check VaR/CVaR on own values
    exp = port.terminal_statistics(v, 100.0)
    for k in ("var_95", "cvar_95", "prob_profit", "expected_vol"):
        assert abs(res[k] - exp[k]) < 1e-9 * max(1.0, abs(exp[k])), k
    assert res["cvar_95"] > res["var_95"] > 0


def test_full_size_c5_materialised_fan_chart():
    """BASELINE config 5 at its full size: 4,194,304 x 1024 fp32 paths stored time-major (17.2 GB); the fan chart is
    np.percentile's, bit for bit, on rows sampled across the horizon (the first step, where a row is one draw; the
    middle; the end), and the terminal statistics are NumPy's on the last row."""
    n, t = 4_194_304, 1024
    res = simulate_gbm(n, t, 100.0, 0.05, 0.2, seed=9, return_paths="device", hist_bins=2048)
    paths = res["paths"]
    assert paths.shape == (t + 1, n) and paths.dtype == torch.float32
    last = paths[-1].cpu().numpy().astype(np.float64)
    for p, v in zip(port.REF_TERMINAL_PCTS, np.percentile(last, port.REF_TERMINAL_PCTS)):
        assert res["percentiles"][f"p{p}"] == v
    assert np.array_equal(res["histogram"]["counts"], np.histogram(last, bins=2048, range=res["histogram"]["range"])[0])
    for row in (0, 1, 2, 17, 511, 800, 1023, 1024):
        ref = np.percentile(paths[row].cpu().numpy().astype(np.float64), port.REF_PATH_PCTS)
        got = [res["path_percentiles"][f"p{p}"][row] for p in port.REF_PATH_PCTS]
        assert list(ref) == got
    assert torch.equal(paths[-1], res["final_prices"])


@pytest.mark.parametrize("mode,dtype", [("garch", "float32"), ("garch", "float64"), ("gbm", "float32")])
def test_fan_chart_by_regeneration_equals_stored_paths(engine, mode, dtype):
    """the fan chart (app.py:643-645) without keeping the paths.  Chunks of the path range
    take turns in one buffer and are simulated again before every stage that reads them; the result must
    be the very same order statistics as from the fully stored matrix."""
    n, t, s0 = 20_011, 37, 100.0
    pool = None
    if mode == "garch":
        pool = port.bootstrap_pool(np.random.RandomState(5).standard_t(4, 1250) * 0.014)
        kw = dict(n_paths=n, n_steps=t, s0=s0, var0=port.initial_variance(0.02 * np.sqrt(252)), dtype=dtype, seed=31,
                  store_paths=True, fan_chart=True, hist_bins=64)
        full, chunked = SimSpec(**kw), SimSpec(fan_chunk_paths=3000, **kw)
    else:
        kw = dict(dtype=dtype, seed=31, store_paths=True, fan_chart=True, hist_bins=64, hist_range=(50.0, 200.0))
        full = SimSpec.gbm(n, t, s0, 0.05, 0.2, 1 / 252, **kw)
        chunked = SimSpec.gbm(n, t, s0, 0.05, 0.2, 1 / 252, fan_chunk_paths=3000, **kw)
    a = engine.simulate(full, pool=pool)
    b = engine.simulate(chunked, pool=pool)
    assert b.paths is None and a.paths is not None
    assert torch.equal(a.terminal, b.terminal)
    for k in a.fan:
        assert np.array_equal(a.fan[k], b.fan[k]), k
    ref_fan = np.percentile(a.paths.cpu().numpy().astype(np.float64), port.REF_PATH_PCTS, axis=1)
    assert np.array_equal(np.stack([b.fan[f"p{p}"] for p in port.REF_PATH_PCTS]), ref_fan)
    assert a.pct == b.pct and a.n == b.n and a.n_above == b.n_above
    assert a.tail_count == b.tail_count and np.array_equal(a.hist, b.hist)
    assert abs(a.mean - b.mean) < 1e-12 * abs(a.mean) and abs(a.std - b.std) < 1e-10 * a.std
    assert abs(a.tail_sum - b.tail_sum) < 1e-12 * abs(a.tail_sum)


@pytest.mark.parametrize("dtype,n,t,tiny", [("float32", 300_007, 120, False), ("float64", 50_003, 64, False),
                                              ("float32", 300_007, 120, True)])
def test_terminal_selection_beside_the_fan_chart_changes_nothing(engine, dtype, n, t, tiny):
    """mc_run enqueues the terminal-value problem of a fan-chart request (one row: brackets, two passes, scans, tail,
    histogram) on a stream of its own, underneath the fan chart's passes over the whole matrix.  Every number must be
    the one the single-stream order gives (mc_debug_set_tuning bit 1), also when tiny side lists force the passes
    to stream the rows again and brackets to be revisited."""
    spec = SimSpec.gbm(n, t, 100.0, 0.05, 0.2, 1 / 252, dtype=dtype, seed=77, store_paths=True, fan_chart=True,
                       hist_bins=512, hist_range=(40.0, 260.0))
    try:
        engine.debug_set_tuning(side_cap=32 if tiny else -1, no_overlap=True)
        a = engine.simulate(spec)
        engine.debug_set_tuning(side_cap=32 if tiny else -1, no_overlap=False)
        for _ in range(3):                      # back to back: nothing of one run may leak into the next
            b = engine.simulate(spec)
            assert a.pct == b.pct and a.pct_bounds == b.pct_bounds
            assert (a.n, a.n_above, a.tail_count) == (b.n, b.n_above, b.tail_count)
            assert (a.mean, a.std, a.tail_sum, a.min, a.max) == (b.mean, b.std, b.tail_sum, b.min, b.max)
            assert np.array_equal(a.hist, b.hist)
            for k in a.fan:
                assert np.array_equal(a.fan[k], b.fan[k]), k
    finally:
        engine.debug_set_tuning()
    last = b.paths[-1].cpu().numpy().astype(np.float64)
    assert [b.pct[f"p{p}"] for p in port.REF_TERMINAL_PCTS] == list(np.percentile(last, port.REF_TERMINAL_PCTS))
    assert np.array_equal(b.hist, np.histogram(last, bins=512, range=(40.0, 260.0))[0])


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_brackets_that_miss_are_revisited_also_with_the_terminal_selection_on_its_own_stream(engine, dtype):
    """Injected shocks SORTED across the paths: every row of the path matrix is sorted, so the pilot (the first 131072
    values of a row) holds the smallest ones and the brackets of the upper percentiles miss.  The selection then narrows
    the gap beside the bracket in further passes -- after the closing stage has already been enqueued once -- while the
    terminal-value problem runs on the handle's own stream.  Results: NumPy's, bit for bit."""
    n, t = 300_007, 3
    z = np.sort(np.random.RandomState(3).normal(size=(t, n)), axis=1)
    spec = SimSpec.gbm(n, t, 100.0, 0.05, 0.2, 1 / 252, dtype=dtype, shock="injected", store_paths=True, fan_chart=True,
                       hist_bins=128, hist_range=(90.0, 112.0))
    out = engine.simulate(spec, shocks=z)
    paths = out.paths.cpu().numpy().astype(np.float64)
    assert np.all(np.diff(paths[-1]) >= 0)                                    # the premise: sorted rows
    assert out.launches > 40                       # 26 launches when the brackets hold; measured here: 130 - 143
    ref_fan = np.percentile(paths, port.REF_PATH_PCTS, axis=1)
    assert np.array_equal(np.stack([out.fan[f"p{p}"] for p in port.REF_PATH_PCTS]), ref_fan)
    assert [out.pct[f"p{p}"] for p in port.REF_TERMINAL_PCTS] == list(np.percentile(paths[-1], port.REF_TERMINAL_PCTS))
    assert np.array_equal(out.hist, np.histogram(paths[-1], bins=128, range=(90.0, 112.0))[0])
    thr = out.pct["p5"]
    assert out.tail_count == int((paths[-1] <= thr).sum())


def test_regeneration_restores_rows_when_the_side_lists_do_not_cover_a_pass(engine):
    """MC_FLAG_PATHS_VOLATILE: from the second narrowing pass on a chunk works from the values pass 1 kept; with
    32-entry lists they overflow, and the handle must simulate the chunk again by itself -- same fan chart."""
    kw = dict(dtype="float32", seed=77, store_paths=True, fan_chart=True)
    n, t = 300_007, 9
    full = SimSpec.gbm(n, t, 100.0, 0.05, 0.2, 1 / 252, **kw)
    a = engine.simulate(full)
    ref_fan = np.percentile(a.paths.cpu().numpy().astype(np.float64), port.REF_PATH_PCTS, axis=1)
    launches = {}
    try:
        for cap in (None, 32):
            engine.debug_set_tuning(side_cap=-1 if cap is None else cap)
            b = engine.simulate(SimSpec.gbm(n, t, 100.0, 0.05, 0.2, 1 / 252, fan_chunk_paths=100_000, **kw))
            assert np.array_equal(np.stack([b.fan[f"p{p}"] for p in port.REF_PATH_PCTS]), ref_fan)
            assert a.pct == b.pct and torch.equal(a.terminal, b.terminal)
            launches[cap] = b.launches
    finally:
        engine.debug_set_tuning()
    assert launches[32] > launches[None]          # the extra simulations


def test_drop_in_without_paths_still_has_the_fan_chart():
    df = _frame()
    kept = garch_monte_carlo(df, 30_000, 25, 100.0, seed=3, dtype="float32", return_paths="numpy")
    import montecarlo_b200.api as api
    old = api.FAN_BUFFER_BYTES
    api.FAN_BUFFER_BYTES = 26 * 4 * 8192          # 8192 paths per chunk
    try:
        lean = garch_monte_carlo(df, 30_000, 25, 100.0, seed=3, dtype="float32", return_paths=False)
    finally:
        api.FAN_BUFFER_BYTES = old
    assert lean["paths"] is None
    for k in ("p5", "p25", "p50", "p75", "p95"):                                   # app.py:972-981
        assert np.array_equal(lean["path_percentiles"][k], kept["path_percentiles"][k])
    assert lean["percentiles"] == kept["percentiles"]
    assert garch_monte_carlo(df, 1000, 5, 100.0, seed=3, return_paths=False, fan_chart=False)["path_percentiles"] is None


def test_concurrent_callers_from_threads():
    """The app runs sessions in threads (app.py:1392): concurrent calls must not interfere."""
    import threading
    df = _frame()
    expect = {s: garch_monte_carlo(df, 4000, 30, 100.0, seed=s, dtype="float32") for s in range(6)}
    got, errs = {}, []

    def work(s):
        try:
            for _ in range(3):
                got[s] = garch_monte_carlo(df, 4000, 30, 100.0, seed=s, dtype="float32")
        except Exception as exc:       # pragma: no cover
            errs.append(exc)

    threads = [threading.Thread(target=work, args=(s,)) for s in range(6)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errs
    for s in range(6):
        assert np.array_equal(got[s]["paths"], expect[s]["paths"])
        assert got[s]["percentiles"] == expect[s]["percentiles"] and got[s]["cvar_95"] == expect[s]["cvar_95"]
