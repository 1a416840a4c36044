"""The oracle against the fixtures generated from the UNMODIFIED reference
(oracle/make_golden.py ran /root/reference/app.py:586-657 in the build container)."""
import os

import numpy as np
import pytest

from oracle import mc_numpy as port
from oracle import philox

SCALARS = ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95")


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def _check(res, g):
    for k in SCALARS:
        assert res[k] == g[k], k                     # bit-exact: same NumPy ops on the same inputs
    got = np.array([res["percentiles"][f"p{p}"] for p in port.REF_TERMINAL_PCTS])
    assert np.array_equal(got, g["percentiles"])
    fan = np.stack([res["path_percentiles"][f"p{p}"] for p in port.REF_PATH_PCTS])
    assert np.array_equal(fan, g["path_percentiles"])


@pytest.mark.parametrize("name", ["ref_small_bootstrap.npz", "ref_ragged_heavytail.npz"])
def test_port_matches_reference_injected_indices(golden_dir, name):
    g = _load(golden_dir, name)
    res = port.garch_monte_carlo_port(g["returns"], float(g["rvol_last"]), int(g["n_sims"]), int(g["n_days"]),
                                      float(g["s0"]), indices=g["indices"])
    assert np.array_equal(res["paths"], g["paths"])
    _check(res, g)


def test_port_matches_reference_normal_shocks(golden_dir):
    g = _load(golden_dir, "ref_normal_through_garch.npz")
    res = port.garch_monte_carlo_port(g["returns"], float(g["rvol_last"]), int(g["n_sims"]), int(g["n_days"]),
                                      float(g["s0"]), shocks=g["shocks"])
    assert np.array_equal(res["paths"], g["paths"])
    _check(res, g)


def test_port_matches_reference_c1_seeded(golden_dir):
    """BASELINE config 1: 10k x 252 under np.random.seed(42); the draws are regenerated
    with the same legacy RandomState (np.random.choice == pool[randint])."""
    g = _load(golden_dir, "ref_c1_seed42.npz")
    res = port.garch_monte_carlo_port(g["returns"], float(g["rvol_last"]), int(g["n_sims"]), int(g["n_days"]),
                                      float(g["s0"]), rng=np.random.RandomState(int(g["np_seed"])))
    assert np.array_equal(res["final_prices"][::97], g["final_every_97"])
    _check(res, g)


def test_choice_is_pool_of_randint():
    pool = np.random.RandomState(1).normal(size=1250)
    a = np.random.RandomState(7).choice(pool, size=1000)
    b = pool[np.random.RandomState(7).randint(0, len(pool), size=1000)]
    assert np.array_equal(a, b)


def test_short_history_returns_none():
    assert port.garch_monte_carlo_port(np.zeros(99), 0.3, 10, 5, 100.0) is None
    assert port.garch_monte_carlo_port(np.random.RandomState(0).normal(size=100), 0.3, 10, 5, 100.0,
                                       rng=np.random.RandomState(0)) is not None


def test_philox_known_answers():
    for ctr, key, expect in philox.KAT_VECTORS:
        out = philox.philox4x32([np.array([c]) for c in ctr], key)
        assert tuple(int(o[0]) for o in out) == expect


def test_philox_normals_are_standard():
    z = philox.normal_shocks(0x1234, 0, 100_000, 8)
    n = z.size
    assert abs(z.mean()) < 4 / np.sqrt(n)
    assert abs(z.var() - 1) < 4 * np.sqrt(2 / n)
    assert abs((z ** 4).mean() - 3) < 0.1
    # consecutive outputs of a pair are uncorrelated
    assert abs(np.mean(z[0] * z[1])) < 4 / np.sqrt(z.shape[1])
    z64 = philox.normal_shocks(0x1234, 0, 1000, 8, dtype=np.float64)
    assert np.abs(z64).max() < 7 and abs(z64.mean()) < 0.05


def test_shard_invariance_of_the_bit_model():
    full = philox.normal_shocks(99, 0, 64, 10)
    part = philox.normal_shocks(99, 40, 24, 10)
    assert np.array_equal(full[:, 40:], part)
    idx = philox.bootstrap_indices(5, 0, 1000, 7, 1250)
    assert idx.min() >= 0 and idx.max() < 1250
    assert np.array_equal(idx[:, 300:], philox.bootstrap_indices(5, 300, 700, 7, 1250))


def test_fp32_bootstrap_index_is_the_exact_floor():
    # the fp32 kernels compute floor(field * H / 2^23) with one round-down FFMA (mc_rng.cuh: boot_index_bits);
    # the model of that instruction must equal the integer floor everywhere, the ties field * H = k 2^23 included
    # (round-to-nearest got half of those one index low, and field 0 onto index -1)
    rs = np.random.RandomState(0)
    for h in (100, 1250, 1051, 4096, 65536):
        f = rs.randint(0, 1 << 23, size=200_000).astype(np.uint32)
        f[:4] = [0, 1, (1 << 23) - 1, 1 << 22]
        f[4:4 + 64] = (np.arange(64, dtype=np.uint32) << 17)           # many ties for the powers of two
        exact = (f.astype(np.int64) * h) >> 23
        got = philox._boot_index32(f, h)
        assert np.array_equal(got, exact)
        assert got.min() >= 0 and got.max() < h
    idx = philox.bootstrap_indices(5, 0, 20_000, 10, 1250, dtype=np.float32)
    assert idx.min() >= 0 and idx.max() < 1250
    counts = np.bincount(idx.ravel(), minlength=1250)
    assert abs(counts.std() / np.sqrt(counts.mean()) - 1) < 0.1       # multinomial spread


def test_gbm_closed_form():
    n, t, s0, mu, sigma, dt = 200_000, 12, 100.0, 0.05, 0.2, 1 / 252
    z = philox.normal_shocks(3, 0, n, t)
    final = port.gbm_paths(z, s0, mu, sigma, dt)[-1]
    cf = port.gbm_closed_form(s0, mu, sigma, t * dt, q=[0.05, 0.5, 0.95])
    assert abs(final.mean() - cf["mean"]) < 4 * cf["std"] / np.sqrt(n)
    assert abs(final.std() / cf["std"] - 1) < 0.02
    assert np.allclose(np.percentile(final, [5, 50, 95]), cf["quantiles"], rtol=2e-3)


def test_portfolio_covariance():
    d, n, t, dt = 4, 100_000, 8, 1 / 252
    corr = np.full((d, d), 0.3) + 0.7 * np.eye(d)
    chol = np.linalg.cholesky(corr)
    sigma = 0.15 + 0.01 * np.arange(d)
    z = philox.normal_shocks(11, 0, n, t, n_assets=d)
    v, s_t = port.portfolio_terminal(z, 100.0, 0.05, sigma, chol, np.full(d, 1 / d), dt)
    x = np.log(s_t / 100.0)
    cov = np.cov(x)
    expect = corr * np.outer(sigma, sigma) * t * dt
    assert np.allclose(cov, expect, rtol=0.05, atol=1e-6)
    assert abs(v.mean() - 100.0 * np.exp(0.05 * t * dt)) < 4 * v.std() / np.sqrt(n)
