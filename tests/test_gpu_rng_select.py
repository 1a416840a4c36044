"""GPU parity, part 1: generator bit-parity, exact selection, histogram (through the C ABI)."""
import numpy as np
import pytest
import torch

from oracle import philox
from montecarlo_b200 import SimSpec

pytestmark = pytest.mark.gpu


def test_philox_known_answers_on_device(engine):
    for ctr, key, expect in philox.KAT_VECTORS:
        out = engine.debug_philox(np.array([ctr], dtype=np.uint32), key[0], key[1])
        assert tuple(int(v) for v in out[0]) == expect


def test_philox_random_counters_match_bit_model(engine):
    rs = np.random.RandomState(0)
    ctr = rs.randint(0, 2 ** 32, size=(4096, 4), dtype=np.uint64).astype(np.uint32)
    out = engine.debug_philox(ctr, 0xDEADBEEF, 0x12345678)
    ref = philox.philox4x32([ctr[:, i] for i in range(4)], (0xDEADBEEF, 0x12345678))
    assert np.array_equal(out, np.stack(ref, axis=1))


@pytest.mark.parametrize("dtype,tol", [("float32", 5e-5), ("float64", 1e-12)])
def test_device_normals_match_bit_model(engine, dtype, tol):
    n, t, seed, off = 3000, 11, 0x1234567890ABCDEF, (1 << 33) + 5     # counter high word exercised
    spec = SimSpec.gbm(n, t, 100.0, 0.05, 0.2, dtype=dtype, seed=seed)
    z = engine.debug_shocks(spec, path_offset=off).cpu().numpy().astype(np.float64)
    ref = philox.normal_shocks(seed, off, n, t, dtype=np.dtype(dtype))
    assert z.shape == ref.shape
    # MUFU lg2/sqrt are approximate: the worst absolute error sits near z = 0 (lg2 of a uniform close
    # to 1); on average the device normal is within 1e-6 of the fp64 bit model
    assert np.abs(z - ref).max() < tol
    assert np.abs(z - ref).mean() < (1e-6 if dtype == 'float32' else 1e-13)
    assert abs((z - ref).mean()) < (1e-7 if dtype == 'float32' else 1e-14)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
@pytest.mark.parametrize("n,t,H", [(2000, 9, 1250), (60_000, 25, 1024), (20_000, 10, 65536), (30_000, 20, 100)])
def test_device_bootstrap_draws_match_bit_model(engine, dtype, n, t, H):
    """Pool lengths with a power of two in them put many 23-bit fields exactly on an index boundary
    (field * H = k 2^23: one draw in 8192 for H = 1024): the fp32 single-FMA floor must take those right."""
    seed = 77
    pool = np.random.RandomState(1).normal(size=H)
    spec = SimSpec(n_paths=n, n_steps=t, s0=100.0, dtype=dtype, seed=seed, var0=4e-4)
    z = engine.debug_shocks(spec, pool=pool).cpu().numpy()
    idx = philox.bootstrap_indices(seed, 0, n, t, H, dtype=np.dtype(dtype))
    assert np.array_equal(z, pool.astype(dtype)[idx])


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("n", [1, 2, 7, 1000, 100_003])
def test_percentiles_exact(engine, dtype, n):
    rs = np.random.RandomState(n)
    x = (100 * np.exp(rs.standard_t(3, n) * 0.2))
    x[rs.randint(0, n, n // 3)] = x[0]                 # ties
    x[rs.randint(0, n, max(1, n // 10))] *= -1         # negatives
    xt = torch.tensor(x, dtype=dtype, device=engine.device)
    pcts = [0, 1, 5, 10, 25, 50, 75, 90, 95, 99, 100]
    got = engine.percentiles(xt, pcts)[0]
    ref = np.percentile(xt.cpu().numpy().astype(np.float64), pcts)
    assert np.array_equal(got, ref)


def test_percentiles_rows_and_degenerate(engine):
    rs = np.random.RandomState(3)
    x = rs.lognormal(size=(37, 5001)).astype(np.float32)
    x[5, :] = 3.25                                      # a constant row (paths[0] is all s0, app.py:613)
    x[6, :2500] = 3e38                              # near-overflow values sort last
    x[7, :100] = np.inf                              # +inf beyond every requested rank
    xt = torch.tensor(x, device=engine.device)
    got = engine.percentiles(xt, [5, 25, 50, 75, 95])
    ref = np.percentile(x.astype(np.float64), [5, 25, 50, 75, 95], axis=1).T
    assert np.array_equal(got, ref)
    # strided rows (a column block of a wider matrix)
    wide = torch.tensor(rs.normal(size=(9, 4000)), device=engine.device)
    view = wide[:, :3000]
    assert np.array_equal(engine.percentiles(view, [50])[:, 0], np.percentile(view.cpu().numpy(), 50, axis=1))


@pytest.mark.parametrize("mode", ["default", "tiny_lists", "off"])
def test_percentiles_side_lists_and_their_fallbacks(engine, mode):
    """fp32 pass 1 keeps the in-bracket values in side lists that the later passes read instead of the rows;
    a list that overflows (forced here with 32-entry lists) or side lists switched off fall back to streaming
    the rows: the result is the exact np.percentile either way."""
    engine.debug_set_tuning(side_cap=32 if mode == "tiny_lists" else -1, no_side_lists=mode == "off")
    try:
        rs = np.random.RandomState(17)
        x = (50 * np.exp(0.3 * rs.standard_normal((11, 600_000)))).astype(np.float32)
        x[3, ::7] = x[3, 0]                                    # heavy ties
        x[4, :] = np.sort(x[4, :])                             # sorted row: the leading pilot sample misses every bracket
        x[5, :300_000] = 1.0                                   # half the row is one value
        xt = torch.tensor(x, device=engine.device)
        for pcts in ([5, 25, 50, 75, 95], [1, 5, 10, 25, 50, 75, 90, 95, 99]):
            ref = np.percentile(x.astype(np.float64), pcts, axis=1).T
            assert np.array_equal(engine.percentiles(xt, pcts), ref)
        one = torch.tensor(x[0], device=engine.device)
        assert np.array_equal(engine.percentiles(one, [1, 50, 99])[0], np.percentile(x[0].astype(np.float64), [1, 50, 99]))
    finally:
        engine.debug_set_tuning()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_histogram_matches_numpy(engine, dtype):
    rs = np.random.RandomState(9)
    n, bins, lo, hi = 200_000, 2048, 40.0, 250.0
    x = 100 * np.exp(rs.normal(size=n) * 0.3)
    edges = np.linspace(lo, hi, bins + 1)
    x[:bins + 1] = edges                                # values exactly on every edge
    x[bins + 1:bins + 399] = np.nextafter(edges[1:399], 0)
    x[3000:3010] = [lo - 1e-9, hi + 1e-9, lo, hi, 0, 1e9, -5, hi, lo, 100.0]
    xt = torch.tensor(x, dtype=dtype, device=engine.device)
    got = engine.histogram(xt, bins, lo, hi)
    ref = np.histogram(xt.cpu().numpy().astype(np.float64), bins=bins, range=(lo, hi))[0]
    assert np.array_equal(got, ref)
    assert engine.histogram(xt[:0], 16, 0.0, 1.0).sum() == 0
