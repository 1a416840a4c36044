"""Randomised sweep of whole jobs through `mc_run`: random mode, precision, sizes (1 path to 300k, 1 step to 700),
percentile lists up to the ABI's limits, histogram sizes, pool lengths.  Every reduction the device returns is
recomputed with NumPy -- the reference's own calls (app.py:636-656) -- on the values the device produced."""
import numpy as np
import pytest
import torch

from montecarlo_b200 import SimSpec
from oracle import mc_numpy as port

pytestmark = pytest.mark.gpu


def _random_job(rs):
    mode = rs.choice(["garch", "gbm", "garch_normal"])
    dtype = str(rs.choice(["float32", "float64"]))
    n = int(rs.choice([1, 2, 5, 31, 1000, 4099, 65_537, 300_000]))
    t = int(rs.choice([1, 2, 5, 7, 64, 253, 700])) if n <= 65_537 else int(rs.choice([1, 5, 33]))
    store = bool(rs.uniform() < 0.6)
    fan = store and bool(rs.uniform() < 0.8)
    n_p = int(rs.randint(1, 17))
    pcts = tuple(sorted(set(np.round(rs.uniform(0, 100, n_p), int(rs.choice([0, 1]))).tolist())))
    path_pcts = tuple(sorted(set(np.round(rs.uniform(0, 100, int(rs.randint(1, 9))), 0).tolist())))
    tail = float(rs.choice(pcts)) if rs.uniform() < 0.7 else None
    bins = int(rs.choice([0, 1, 7, 64, 2048, 65536]))
    hist_range = None if rs.uniform() < 0.4 else (float(rs.uniform(20, 90)), float(rs.uniform(110, 400)))
    kw = dict(dtype=dtype, seed=int(rs.randint(1, 1 << 30)), store_paths=store, fan_chart=fan, pcts=pcts,
              path_pcts=path_pcts, tail_pct=tail, hist_bins=bins, hist_range=hist_range if bins else None)
    pool = None
    if mode == "gbm":
        spec = SimSpec.gbm(n, t, 100.0, float(rs.uniform(-0.2, 0.3)), float(rs.uniform(0.05, 0.6)), 1 / 252, **kw)
    else:
        h = int(rs.choice([100, 128, 777, 1024, 1250, 4096, 20_000]))
        pool = port.bootstrap_pool(rs.standard_t(4, h) * 0.014)
        spec = SimSpec(n_paths=n, n_steps=t, s0=100.0, var0=float(rs.uniform(5e-5, 6e-4)),
                       shock="normal" if mode == "garch_normal" else "bootstrap", **kw)
    return spec, pool


@pytest.mark.parametrize("seed", range(40))
def test_random_job(engine, seed):
    rs = np.random.RandomState(5000 + seed)
    spec, pool = _random_job(rs)
    try:
        out = engine.simulate(spec, pool=pool)
    except RuntimeError as exc:
        raise AssertionError(f"{exc} -- {spec} pool {None if pool is None else len(pool)}") from exc
    final = out.terminal.cpu().numpy().astype(np.float64)
    s0, n = spec.s0, spec.n_paths
    assert final.shape == (n,) and np.isfinite(final).all() and (final > 0).all()
    assert out.n == n and out.n_above == int((final > s0).sum())
    assert abs(out.mean - final.mean()) <= 1e-12 * abs(final.mean())
    assert abs(out.std - final.std()) <= 1e-9 * final.std() + 1e-12 * abs(final.mean())
    assert out.min == final.min() and out.max == final.max()
    for p in spec.pcts:
        assert out.pct[f"p{p:g}"] == np.percentile(final, p), p
    if spec.tail_pct is not None:
        thr = np.percentile(final, spec.tail_pct)
        sel = final[final <= thr]                                        # app.py:656
        assert out.tail_count == sel.size and abs(out.tail_sum - sel.sum()) <= 1e-12 * abs(sel.sum())
    if spec.hist_bins:
        lo, hi = spec.hist_range if spec.hist_range else (final.min(), final.max())
        if lo < hi:
            counts, _ = np.histogram(final, bins=spec.hist_bins, range=(lo, hi))
            assert np.array_equal(out.hist, counts)
    if spec.store_paths:
        paths = out.paths.cpu().numpy().astype(np.float64)
        assert paths.shape == (spec.n_steps + 1, n) and np.all(paths[0] == s0) and np.array_equal(paths[-1], final)
        if spec.fan_chart:
            ref = np.percentile(paths, spec.path_pcts, axis=1)
            assert np.array_equal(np.stack([out.fan[f"p{p:g}"] for p in spec.path_pcts]), ref)
    else:
        assert out.paths is None


@pytest.mark.parametrize("seed", range(12))
def test_random_fan_chart_by_regeneration(engine, seed):
    """Random chunkings of a stored-paths job (every chunk a virtual rank of the stage protocol, simulated again
    only when a pass needs its rows): the same exact results as the job kept whole."""
    rs = np.random.RandomState(7000 + seed)
    n = int(rs.choice([1000, 4099, 50_001, 120_000]))
    t = int(rs.choice([1, 3, 20, 61]))
    dtype = str(rs.choice(["float32", "float64"]))
    chunk = int(rs.choice([max(4096, n // 7), max(4096, n // 3), 4096]))
    pcts = tuple(sorted(set(np.round(rs.uniform(0, 100, int(rs.randint(1, 10))), 0).tolist())))
    kw = dict(dtype=dtype, seed=int(rs.randint(1, 1 << 30)), store_paths=True, fan_chart=True, pcts=pcts,
              tail_pct=float(pcts[0]), hist_bins=int(rs.choice([0, 64])), hist_range=(50.0, 200.0))
    pool = None
    if rs.uniform() < 0.5:
        mk = lambda **e: SimSpec.gbm(n, t, 100.0, 0.05, 0.25, 1 / 252, **kw, **e)
    else:
        pool = port.bootstrap_pool(rs.standard_t(4, int(rs.choice([128, 1250]))) * 0.014)
        mk = lambda **e: SimSpec(n_paths=n, n_steps=t, s0=100.0, var0=2e-4, **kw, **e)
    whole = engine.simulate(mk(), pool=pool)
    parts = engine.simulate(mk(fan_chunk_paths=chunk), pool=pool)
    assert torch.equal(whole.terminal, parts.terminal)
    assert whole.pct == parts.pct and whole.n_above == parts.n_above and whole.tail_count == parts.tail_count
    for k in whole.fan:
        assert np.array_equal(whole.fan[k], parts.fan[k]), k
    if whole.hist is not None:
        assert np.array_equal(whole.hist, parts.hist)
    assert abs(whole.mean - parts.mean) <= 1e-12 * abs(whole.mean)
    assert abs(whole.tail_sum - parts.tail_sum) <= 1e-12 * abs(whole.tail_sum)
