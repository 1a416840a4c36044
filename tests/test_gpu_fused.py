"""GPU parity, part 4: the fused closing (csrc/mc_fused.cu) -- the reductions of app.py:636-656 for runs without a
fan chart in four launches, and path-sharded runs whose two exchanges the kernels carry out through peer memory.
Everything is held to NumPy on the device's own terminal values: percentiles, histogram, counts exact."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import mc_numpy as port
from montecarlo_b200 import PortfolioSpec, SimSpec
from montecarlo_b200 import _lib as L
from montecarlo_b200.spec import make_params, make_portfolio, shard_range

pytestmark = pytest.mark.gpu
PCTS = port.REF_TERMINAL_PCTS


def _check(out, spec, s0):
    v = out.terminal.cpu().numpy().astype(np.float64)
    exp = port.terminal_statistics(v, s0, spec.pcts)
    for p in spec.pcts:
        assert out.pct[f"p{p:g}"] == exp["percentiles"][f"p{p}"], p
    assert out.n == v.size and out.n_above == int((v > s0).sum())
    assert out.min == v.min() and out.max == v.max()
    summ = out.summary()
    assert abs(out.mean - v.mean()) <= 1e-12 * abs(v.mean())
    assert abs(out.std - v.std()) <= 1e-9 * v.std() + 1e-10 * abs(v.mean())     # sqrt of a cancellation when all values agree
    if spec.tail_pct is not None and spec.pcts:
        assert summ["var_95"] == exp["var_95"]
        thr = np.percentile(v, spec.tail_pct)
        assert out.tail_count == int((v <= thr).sum())
        assert abs(out.tail_sum - v[v <= thr].sum()) <= 1e-12 * abs(v[v <= thr].sum())
    if spec.hist_bins:
        counts, _ = np.histogram(v, bins=spec.hist_bins, range=spec.hist_range)
        assert np.array_equal(out.hist, counts)
    return v


def _pool():
    return port.bootstrap_pool(np.random.RandomState(5).standard_t(4, 1250) * 0.014)


CASES = {
    "gbm32": lambda n: (SimSpec.gbm(n, 64, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=11, hist_bins=2048,
                                    hist_range=(40.0, 250.0)), None),
    "gbm32_manybins": lambda n: (SimSpec.gbm(n, 64, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=12, hist_bins=5000,
                                             hist_range=(40.0, 250.0)), None),
    "gbm32_nohist": lambda n: (SimSpec.gbm(n, 37, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=13), None),
    "garch32": lambda n: (SimSpec(n_paths=n, n_steps=60, s0=100.0, var0=1.6e-4, dtype="float32", seed=14, hist_bins=512,
                                  hist_range=(20.0, 400.0)), _pool()),
    "garch64": lambda n: (SimSpec(n_paths=n, n_steps=60, s0=100.0, var0=1.6e-4, dtype="float64", seed=15, hist_bins=512,
                                  hist_range=(20.0, 400.0)), _pool()),
    "gbm64": lambda n: (SimSpec.gbm(n, 30, 100.0, 0.05, 0.2, 1 / 252, dtype="float64", seed=16, hist_bins=100,
                                    hist_range=(50.0, 200.0)), None),
    "gbm32_stored": lambda n: (SimSpec.gbm(n, 16, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=17, store_paths=True,
                                           hist_bins=64, hist_range=(50.0, 200.0)), None),
}


@pytest.mark.parametrize("n", [1, 2, 37, 5000, 300_001, 2_000_003])
@pytest.mark.parametrize("case", sorted(CASES))
def test_fused_closing_matches_numpy(engine, case, n):
    spec, pool = CASES[case](n)
    before = engine.debug_counter(0), engine.debug_counter(1)
    out = engine.simulate(spec, pool=pool)
    _check(out, spec, 100.0)
    assert engine.debug_counter(0) == before[0] + 1          # the fused closing ran ...
    assert engine.debug_counter(1) == before[1]              # ... and stood
    assert out.launches <= 6


def test_fused_closing_portfolio(engine):
    d = 16
    pf = PortfolioSpec(s0=[100.0] * d, mu=[0.05] * d, sigma=[0.15 + 0.01 * i for i in range(d)], weights=[1.0 / d] * d,
                       corr=0.3 * np.ones((d, d)) + 0.7 * np.eye(d))
    for dtype in ("float32", "float64"):
        spec = SimSpec(n_paths=200_003, n_steps=21, s0=100.0, shock="normal", var_model="const", dtype=dtype, seed=3,
                       portfolio=pf, hist_bins=256, hist_range=(60.0, 160.0))
        f0 = engine.debug_counter(1)
        out = engine.simulate(spec)
        _check(out, spec, 100.0)
        assert engine.debug_counter(1) == f0


@pytest.mark.parametrize("pcts", [(0, 100), (50,), (0.001, 99.999), tuple(range(1, 17)), ()])
def test_fused_closing_other_percentile_lists(engine, pcts):
    spec = SimSpec.gbm(123_457, 20, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=5, pcts=tuple(float(p) for p in pcts),
                       tail_pct=None)
    out = engine.simulate(spec)
    v = out.terminal.cpu().numpy().astype(np.float64)
    for p in pcts:
        assert out.pct[f"p{float(p):g}"] == np.percentile(v, p), p
    assert out.n == v.size and abs(out.mean - v.mean()) < 1e-12 * v.mean()


def test_fused_closing_degenerate_distributions(engine):
    """All paths equal (zero variance): one histogram bin holds everything, the selection narrows onto a single key.
    And shocks five times wider than the model says: the key window misses, the end bins overflow with candidates."""
    spec = SimSpec.gbm(100_003, 10, 100.0, 0.05, 0.0, 1 / 252, dtype="float32", seed=1)
    out = engine.simulate(spec)
    v = _check(out, spec, 100.0)
    assert v.min() == v.max()
    for dtype in ("float32", "float64"):
        for n, falls_back in ((200_000, False), (3_000_000, True)):
            z = (5.0 * np.random.RandomState(n % 97).standard_normal((8, n))).astype(dtype)
            spec = SimSpec.gbm(n, 8, 100.0, 0.0, 0.4, 1 / 252, dtype=dtype, shock="injected", hist_bins=128,
                               hist_range=(10.0, 1000.0))
            f0 = engine.debug_counter(1)
            out = engine.simulate(spec, shocks=z)
            _check(out, spec, 100.0)
            assert (engine.debug_counter(1) > f0) == falls_back, (dtype, n)


def test_terminal_values_with_ties_and_special_values(engine):
    """Mode R with a three-value pool: a few thousand distinct terminal values, heavy ties in every bin."""
    pool = np.array([-1.0, 0.0, 1.0] * 40)
    for dtype in ("float32", "float64"):
        spec = SimSpec(n_paths=400_001, n_steps=12, s0=100.0, var0=1e-4, omega=0.0, alpha=0.0, beta=1.0, dtype=dtype, seed=2,
                       hist_bins=50, hist_range=(90.0, 110.0))
        out = engine.simulate(spec, pool=pool)
        _check(out, spec, 100.0)


# ---------------------------------------------------------------- sharded runs: ranks emulated on one GPU
def _sharded_emulated(spec, pool, world, pf=None):
    """`world` handles on ONE device, each with its own stream, attached to one another's exchange buffers by plain
    device pointers; every rank enqueues mc_run_sharded, then all are fetched.  The kernels wait for each other's
    flags exactly as they do across NVLink."""
    from montecarlo_b200.engine import Engine, _CudaStageBackend
    engines = [Engine(0) for _ in range(world)]
    dt = torch.float64 if spec.dtype == "float64" else torch.float32
    ptrs = (C.c_void_p * world)()
    for r, e in enumerate(engines):
        local, nbytes = C.c_void_p(), C.c_uint64()
        e.check(e.lib.mc_peer_export(e.handle, world, None, C.byref(local), C.byref(nbytes)))
        ptrs[r] = local.value
    for r, e in enumerate(engines):
        e.check(e.lib.mc_peer_attach(e.handle, r, world, None, ptrs))
    streams = [torch.cuda.Stream(device=0) for _ in range(world)]
    terms, keep = [], []
    ptr = _CudaStageBackend._ptr
    for r, e in enumerate(engines):                 # every buffer first: nothing below may wait for the device
        off, cnt = shard_range(spec.n_paths, r, world)
        pool_t = e.to_device(pool, dt) if pool is not None else None
        if pool_t is not None:
            spec.pool_len = pool_t.numel()
        params = make_params(spec, path_offset=off, n_local=cnt, n_ranks=world)
        terms.append(torch.empty(cnt, dtype=dt, device=e.device))
        keep.append((params, make_portfolio(spec), pool_t))
    torch.cuda.synchronize()
    for r, e in enumerate(engines):                 # rank r's kernels wait (on the device) for the ranks enqueued later
        params, pfs, pool_t = keep[r]
        e.check(e.lib.mc_run_sharded(e.handle, C.byref(params), C.byref(pfs) if pfs else None, ptr(pool_t), None, None,
                                     ptr(terms[r]), C.c_void_p(streams[r].cuda_stream)))
    results = []
    for r, e in enumerate(engines):
        res = L.McResult()
        hist = np.zeros(max(1, spec.hist_bins), dtype=np.uint64)
        e.check(e.lib.mc_fetch_result(e.handle, C.byref(res), C.c_void_p(hist.ctypes.data), None,
                                      C.c_void_p(streams[r].cuda_stream)))
        results.append((res, hist[:spec.hist_bins]))
    torch.cuda.synchronize()
    for e in engines:
        e.close()
    return results, torch.cat(terms)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("case", ["gbm32", "garch32", "garch64", "gbm32_manybins"])
def test_sharded_run_through_peer_memory_equals_one_gpu(engine, case, world):
    n = 400_003
    spec, pool = CASES[case](n)
    single = engine.simulate(spec, pool=pool)
    spec2, _ = CASES[case](n)
    results, terminal = _sharded_emulated(spec2, pool, world)
    assert torch.equal(terminal, single.terminal)
    for res, hist in results:                        # every rank holds the same, complete result
        assert res.status == 0
        assert res.n == n and res.n_above == single.n_above
        assert [res.pct[i] for i in range(len(PCTS))] == list(single.pct.values())
        assert res.tail_count == single.tail_count
        assert abs(res.tail_sum - single.tail_sum) <= 1e-12 * abs(single.tail_sum)
        assert abs(res.mean - single.mean) <= 1e-13 * abs(single.mean)
        assert abs(res.std - single.std) <= 1e-10 * single.std
        assert res.min == single.min and res.max == single.max
        assert np.array_equal(hist.astype(np.int64), single.hist)
    first = results[0][0]
    for res, _ in results[1:]:                       # bit for bit, moments included: same numbers added in rank order
        assert res.mean == first.mean and res.std == first.std and res.tail_sum == first.tail_sum


@pytest.mark.parametrize("dtype,per_rank,steps", [("float32", 8_500_003, 8), ("float64", 6_000_001, 4),
                                                   ("float32", 17_000_001, 4)])
def test_sharded_run_with_bins_too_full_for_whole_keys_in_shared_memory(engine, dtype, per_rank, steps):
    """Eight ranks at the bench's scale: the key-histogram bin of a central percentile then holds 40,000 - 90,000 values.
    fp32, <= 81,920: the select kernel keeps them as 16-bit offsets from the bin's first key; beyond that, and for
    fp64 keys beyond 20,480, it sweeps the ranks' lists (over NVLink on real GPUs) once per narrowing step.  Either way
    the order statistics are NumPy's on the union of the ranks' values, and every rank holds them."""
    world = 8
    n = world * per_rank
    mk = lambda: SimSpec.gbm(n, steps, 100.0, 0.05, 0.2, 1 / 252, dtype=dtype, seed=2024, hist_bins=1024,
                             hist_range=(80.0, 125.0))
    results, terminal = _sharded_emulated(mk(), None, world)
    v = terminal.cpu().numpy().astype(np.float64)
    del terminal
    want = list(np.percentile(v, PCTS))
    counts = np.histogram(v, bins=1024, range=(80.0, 125.0))[0]
    thr = want[list(PCTS).index(5)]
    tail = v[v <= thr]
    first = results[0][0]
    for res, hist in results:
        assert res.status == 0 and res.n == n
        assert [res.pct[i] for i in range(len(PCTS))] == want
        assert np.array_equal(hist.astype(np.int64), counts)
        assert res.tail_count == tail.size and abs(res.tail_sum - tail.sum()) <= 1e-9 * tail.sum()
        assert (res.mean, res.std, res.tail_sum) == (first.mean, first.std, first.tail_sum)
    assert abs(first.mean - v.mean()) <= 1e-9 * v.mean()
    single = engine.simulate(mk())                          # the same job on one handle: its own (local) lists, same code
    assert list(single.pct.values()) == want and single.tail_count == tail.size
    assert np.array_equal(single.hist, counts)


def test_sharded_run_more_ranks_than_paths(engine):
    spec = SimSpec.gbm(5, 9, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=4)
    results, terminal = _sharded_emulated(spec, None, 8)
    v = terminal.cpu().numpy().astype(np.float64)
    for res, _ in results:
        assert res.status == 0 and res.n == 5
        assert [res.pct[i] for i in range(len(PCTS))] == list(np.percentile(v, PCTS))


def test_terminal_given_runs_the_staged_selection_without_a_simulation(engine):
    """MC_FLAG_TERMINAL_GIVEN: what a run falls back to when the fused closing gives up -- here asked for directly."""
    from montecarlo_b200.engine import _CudaStageBackend
    rs = np.random.RandomState(8)
    x = (100 * np.exp(0.3 * rs.standard_normal(250_001))).astype(np.float32)
    term = torch.tensor(x, device=engine.device)
    spec = SimSpec.gbm(x.size, 5, 100.0, 0.0, 0.2, 1 / 252, dtype="float32", hist_bins=77, hist_range=(30.0, 300.0))
    params = make_params(spec)
    params.flags |= L.FLAG_TERMINAL_GIVEN
    res, hist = L.McResult(), np.zeros(77, dtype=np.uint64)
    engine.check(engine.lib.mc_run(engine.handle, C.byref(params), None, None, None, None, _CudaStageBackend._ptr(term),
                                   C.byref(res), C.c_void_p(hist.ctypes.data), None, engine.stream_ptr()))
    v = x.astype(np.float64)
    assert [res.pct[i] for i in range(len(PCTS))] == list(np.percentile(v, PCTS))
    assert res.n == x.size and res.min == v.min() and res.max == v.max()
    assert abs(res.mean - v.mean()) < 1e-12 * v.mean()
    assert np.array_equal(hist.astype(np.int64), np.histogram(v, bins=77, range=(30.0, 300.0))[0])
    assert torch.equal(term, torch.tensor(x, device=engine.device))          # untouched


# ---------------------------------------------------------------- the reference's own call sizes: two launches
@pytest.mark.parametrize("dtype", ["float64", "float32"])
@pytest.mark.parametrize("n,t", [(1, 1), (2, 3), (3, 5), (1000, 20), (5000, 60), (10_000, 252), (16_384, 7), (777, 1)])
def test_small_jobs_take_two_launches_and_equal_the_general_path(engine, dtype, n, t):
    """app.py:1189 / 1195 (5..60 days, 500..5000 simulations) and BASELINE configs[0]: simulation kernel + one block per
    row sorting in shared memory (csrc/mc_small.cu).  Asking for a histogram as well sends the same job down the
    general staged path: order statistics, fan chart, counts, extremes AND mean / std must be bit-identical."""
    pool = _pool()
    kw = dict(n_paths=n, n_steps=t, s0=100.0, var0=port.initial_variance(0.02 * np.sqrt(252)), dtype=dtype, seed=77 + n,
              store_paths=True, fan_chart=True)
    small = engine.simulate(SimSpec(**kw), pool=pool)
    general = engine.simulate(SimSpec(hist_bins=8, hist_range=(1.0, 1e4), **kw), pool=pool)
    assert small.launches == 2 and general.launches > 10
    assert torch.equal(small.paths, general.paths) and torch.equal(small.terminal, general.terminal)
    assert small.pct == general.pct and small.pct_bounds == general.pct_bounds
    for k in general.fan:
        assert np.array_equal(small.fan[k], general.fan[k]), k
    assert (small.n, small.n_above, small.min, small.max) == (general.n, general.n_above, general.min, general.max)
    assert small.mean == general.mean and small.std == general.std
    assert small.tail_count == general.tail_count
    assert abs(small.tail_sum - general.tail_sum) <= 1e-13 * abs(general.tail_sum)
    paths = small.paths.cpu().numpy().astype(np.float64)
    assert np.array_equal(np.stack([small.fan[f"p{p}"] for p in port.REF_PATH_PCTS]),
                          np.percentile(paths, port.REF_PATH_PCTS, axis=1))
    exp = port.terminal_statistics(paths[-1], 100.0, PCTS)
    assert list(small.pct.values()) == [exp["percentiles"][f"p{p}"] for p in PCTS]


def test_small_path_other_models(engine):
    """GBM with device normals and injected shocks, and the largest fp32 row (32768 paths)."""
    for spec, shocks in ((SimSpec.gbm(32_768, 12, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=9, store_paths=True, fan_chart=True), None),
                         (SimSpec.gbm(4_001, 9, 100.0, 0.05, 0.2, 1 / 252, dtype="float64", shock="injected", store_paths=True,
                                      fan_chart=True), np.random.RandomState(1).standard_normal((9, 4_001)))):
        out = engine.simulate(spec, shocks=shocks)
        assert out.launches == 2
        paths = out.paths.cpu().numpy().astype(np.float64)
        assert np.array_equal(np.stack([out.fan[f"p{p}"] for p in port.REF_PATH_PCTS]),
                              np.percentile(paths, port.REF_PATH_PCTS, axis=1))
        exp = port.terminal_statistics(paths[-1], 100.0, PCTS)
        assert list(out.pct.values()) == [exp["percentiles"][f"p{p}"] for p in PCTS]
        s = out.summary()
        for k in ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95"):
            assert abs(s[k] - exp[k]) <= 1e-9 * max(1.0, abs(exp[k])), k


def test_small_path_rows_with_heavy_ties(engine):
    """Three-value pool, constant variance: a handful of distinct prices per row, on both sides of the sort / narrowing split."""
    pool = np.array([-1.0, 0.0, 1.0] * 40)
    for n in (1500, 9000):
        for dtype in ("float32", "float64"):
            spec = SimSpec(n_paths=n, n_steps=6, s0=100.0, var0=1e-4, omega=0.0, alpha=0.0, beta=1.0, dtype=dtype, seed=n,
                           store_paths=True, fan_chart=True)
            out = engine.simulate(spec, pool=pool)
            assert out.launches == 2
            paths = out.paths.cpu().numpy().astype(np.float64)
            assert np.array_equal(np.stack([out.fan[f"p{p}"] for p in port.REF_PATH_PCTS]),
                                  np.percentile(paths, port.REF_PATH_PCTS, axis=1))
            exp = port.terminal_statistics(paths[-1], 100.0, PCTS)
            assert list(out.pct.values()) == [exp["percentiles"][f"p{p}"] for p in PCTS]
            thr = np.percentile(paths[-1], 5.0)
            assert out.tail_count == int((paths[-1] <= thr).sum())


# ---------------------------------------------------------------- jobs in flight on several handles
def test_pipeline_keeps_two_jobs_in_flight_and_returns_the_synchronous_results(engine):
    """`Engine.submit` / `collect` are `mc_run_sharded` + `mc_fetch_result` with one rank: the enqueue and the read of
    `mc_run` as two calls.  `Pipeline(2)` alternates between two handles on two streams; every result must be the one
    the synchronous call gives, bit for bit -- including a job whose key window misses and is reduced again by the
    staged selection inside `collect`."""
    from montecarlo_b200.engine import Pipeline
    pool = _pool()
    jobs = []
    for seed in range(3):
        jobs.append((SimSpec.gbm(300_007 + seed, 30, 100.0, 0.05, 0.2, 1 / 252, seed=seed, hist_bins=256,
                                 hist_range=(60.0, 160.0)), None, None))
        jobs.append((SimSpec(n_paths=200_003, n_steps=25, s0=100.0, var0=2e-4, seed=10 + seed, dtype="float64",
                             hist_bins=64, hist_range=(20.0, 400.0)), pool, None))
    z = (5.0 * np.random.RandomState(3).standard_normal((8, 3_000_000))).astype("float32")       # the window misses
    jobs.append((SimSpec.gbm(3_000_000, 8, 100.0, 0.0, 0.4, 1 / 252, dtype="float32", shock="injected", hist_bins=128,
                             hist_range=(10.0, 1000.0)), None, z))
    pipe = Pipeline(2)
    try:
        got = []
        for spec, pl, sh in jobs:
            if len(pipe._queue) == 2:
                got.append(pipe.collect())
            pipe.submit(spec, pool=pl, shocks=sh)
        while pipe._queue:
            got.append(pipe.collect())
        with pytest.raises(ValueError):
            pipe.submit(SimSpec.gbm(1000, 5, 100.0, 0.05, 0.2, 1 / 252, store_paths=True, fan_chart=True))
        assert pipe.engines[0].debug_counter(1) + pipe.engines[1].debug_counter(1) == 1          # the one fall-back
    finally:
        pipe.close()
    assert len(got) == len(jobs)
    for (spec, pl, sh), out in zip(jobs, got):
        ref = engine.simulate(spec, pool=pl, shocks=sh)
        assert torch.equal(out.terminal, ref.terminal)
        assert out.pct == ref.pct and out.pct_bounds == ref.pct_bounds
        assert (out.n, out.n_above, out.tail_count) == (ref.n, ref.n_above, ref.tail_count)
        assert (out.mean, out.std, out.tail_sum, out.min, out.max) == (ref.mean, ref.std, ref.tail_sum, ref.min, ref.max)
        assert np.array_equal(out.hist, ref.hist)
