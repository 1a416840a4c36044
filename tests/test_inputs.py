"""the routine's inputs from raw close prices.

CPU: the oracle's port of the producers against the fixture written from the UNMODIFIED reference
(`add_all_features`, app.py:210-395; oracle/make_golden.py: inputs_case).  GPU: `mc_prepare_inputs` through the
C ABI against the same fixture and against the port on ragged / first_row variants."""
import os

import numpy as np
import pytest

from oracle import mc_numpy as port
from oracle import ref_loader

# fp64 on both sides, but sums in a different order (block tree vs NumPy pairwise / pandas online window)
POOL_RTOL, STD_RTOL, VAR0_RTOL = 1e-12, 1e-13, 1e-10


@pytest.fixture(scope="module")
def fixture(golden_dir):
    return np.load(os.path.join(golden_dir, "ref_inputs_from_close.npz"))


def test_port_reproduces_the_reference_producers(fixture):
    g = fixture
    for b in range(len(g["lengths"])):
        mine = port.history_inputs(g["close"][b, :g["lengths"][b]], int(g["first_row"]))
        n = int(g["pool_len"][b])
        if n == 0:
            assert mine is None
            continue
        assert np.array_equal(mine["pool"], g["pool"][b, :n])
        assert mine["var0"] == g["var0"][b] and mine["s0"] == g["s0"][b] and mine["pool_std"] == g["pool_std"][b]


def test_port_rows_before_the_first_window_always_go():
    close = ref_loader.synthetic_ohlcv(400, 3)["Close"].values
    a, b = port.history_inputs(close, 0), port.history_inputs(close, 20)
    assert len(a["pool"]) == 380 and np.array_equal(a["pool"], b["pool"])


def _check_series(pool_row, info, exp):
    n = len(exp["pool"])
    assert int(info["pool_len"]) == n and int(info["flags"]) == 0
    assert float(info["s0"]) == exp["s0"]
    assert abs(float(info["pool_std"]) - exp["pool_std"]) <= STD_RTOL * exp["pool_std"]
    assert abs(float(info["var0"]) - exp["var0"]) <= VAR0_RTOL * exp["var0"]
    got = pool_row[:n]
    scale = np.maximum(np.abs(exp["pool"]), 1e-3)           # a return near zero: absolute error of the log
    assert np.max(np.abs(got - exp["pool"]) / scale) <= POOL_RTOL


@pytest.mark.gpu
def test_cuda_producers_against_the_reference_fixture(engine, fixture):
    g = fixture
    pool, info = engine.prepare_inputs(g["close"], g["lengths"], int(g["first_row"]), "float64")
    pool = pool.cpu().numpy()
    for b in range(len(g["lengths"])):
        n = int(g["pool_len"][b])
        if n == 0:
            assert int(info["pool_len"][b]) == 0
            continue
        _check_series(pool[b], info[b], dict(pool=g["pool"][b, :n], s0=g["s0"][b], pool_std=g["pool_std"][b],
                                             var0=g["var0"][b]))
    pool32, info32 = engine.prepare_inputs(g["close"], g["lengths"], int(g["first_row"]), "float32")
    assert np.array_equal(info32, info)
    n0 = int(g["pool_len"][0])
    ref32 = g["pool"][0, :n0].astype(np.float32)
    assert np.max(np.abs(pool32.cpu().numpy()[0, :n0] - ref32) / np.maximum(np.abs(ref32), 1e-3)) <= 2.0 ** -23


@pytest.mark.gpu
@pytest.mark.parametrize("first_row", [0, 20, 57, 199, 700])
def test_cuda_producers_ragged_batch_against_the_port(engine, first_row):
    lengths = np.array([1260, 21, 20, 1, 333, 702, 1259, 800], dtype=np.uint32)
    close = np.full((len(lengths), 1260), 7.0)
    for b, n in enumerate(lengths):
        close[b, :n] = ref_loader.synthetic_ohlcv(int(n), 100 + b, 20.0 + b, heavy_tail=bool(b % 2))["Close"].values
    pool, info = engine.prepare_inputs(close, lengths, first_row, "float64")
    pool = pool.cpu().numpy()
    for b, n in enumerate(lengths):
        exp = port.history_inputs(close[b, :n], first_row)
        if exp is None:
            assert int(info["pool_len"][b]) == 0
        else:
            _check_series(pool[b], info[b], exp)


@pytest.mark.gpu
def test_from_close_equals_the_frame_call(fixture):
    """Same history through pandas + garch_monte_carlo and through the device producers: same seed, same
    bootstrap indices, pools equal to 1e-12 -> results equal to rounding."""
    import pandas as pd
    from montecarlo_b200 import garch_monte_carlo, garch_monte_carlo_from_close
    g = fixture
    res = garch_monte_carlo_from_close(g["close"], 2000, 30, lengths=g["lengths"], seed=77)
    assert [r is None for r in res] == [False, False, False, True, True, False]     # app.py:594-595
    for b in (0, 1, 2, 5):
        n = int(g["lengths"][b])
        inp = port.history_inputs(g["close"][b, :n])
        df = pd.DataFrame({"log_ret": inp["returns"], "rvol_20": inp["rvol_last"]})
        exp = garch_monte_carlo(df, 2000, 30, inp["s0"], seed=77 + b)
        assert np.allclose(res[b]["paths"], exp["paths"], rtol=1e-9, atol=0)
        for k in ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95"):
            assert abs(res[b][k] - exp[k]) <= 1e-7 * max(1.0, abs(exp[k])), k
        for k in exp["path_percentiles"]:
            assert np.allclose(res[b]["path_percentiles"][k], exp["path_percentiles"][k], rtol=1e-9)
    one = garch_monte_carlo_from_close(g["close"][0], 2000, 30, seed=77)
    assert np.array_equal(one["paths"], res[0]["paths"])


@pytest.mark.gpu
def test_from_close_rejects_bad_prices():
    from montecarlo_b200 import garch_monte_carlo_from_close
    close = ref_loader.synthetic_ohlcv(600, 5)["Close"].values.copy()
    close[400] = float("nan")
    with pytest.raises(ValueError):
        garch_monte_carlo_from_close(close, 100, 5, seed=1)


@pytest.mark.gpu
@pytest.mark.parametrize("n_sims,n_days", [(1000, 20), (777, 33), (5000, 60)])
def test_batched_run_is_bit_identical_to_one_run_per_history(fixture, n_sims, n_days):
    """mc_run_batch: one simulation launch over (paths x histories) -- same paths, percentiles and fan chart
    as one mc_run per history with seed + b; moments agree to rounding (different summation order)."""
    from montecarlo_b200 import garch_monte_carlo_from_close
    g = fixture
    loop = garch_monte_carlo_from_close(g["close"], n_sims, n_days, lengths=g["lengths"], seed=5, batched=False)
    batch = garch_monte_carlo_from_close(g["close"], n_sims, n_days, lengths=g["lengths"], seed=5, batched=True)
    assert [r is None for r in batch] == [r is None for r in loop] == [False, False, False, True, True, False]
    for a, b in zip(loop, batch):
        if a is None:
            continue
        assert np.array_equal(a["paths"], b["paths"]) and np.array_equal(a["final_prices"], b["final_prices"])
        assert a["percentiles"] == b["percentiles"]
        for k in a["path_percentiles"]:
            assert np.array_equal(a["path_percentiles"][k], b["path_percentiles"][k])
        assert a["prob_profit"] == b["prob_profit"]
        for k in ("expected_return", "expected_vol", "var_95", "cvar_95"):
            assert abs(a[k] - b[k]) <= 1e-10 * max(1.0, abs(a[k])), k
        exp = port.summarize(b["paths"], float(b["paths"][0, 0]))      # the reference's reductions on these paths
        for k in ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95"):
            assert abs(b[k] - exp[k]) <= 1e-9 * max(1.0, abs(exp[k])), k


@pytest.mark.gpu
def test_batched_run_many_histories_without_paths():
    from montecarlo_b200 import garch_monte_carlo_from_close
    close = np.stack([ref_loader.synthetic_ohlcv(700, 300 + b, 10.0 + b)["Close"].values for b in range(64)])
    res = garch_monte_carlo_from_close(close, 4096, 20, seed=9, return_paths=False, fan_chart=False)
    assert len(res) == 64 and all(r is not None and r["paths"] is None for r in res)
    one = garch_monte_carlo_from_close(close[17], 4096, 20, seed=9 + 17, return_paths=False, fan_chart=False)
    assert one["percentiles"] == res[17]["percentiles"]
    assert np.array_equal(one["final_prices"].cpu().numpy(), res[17]["final_prices"].cpu().numpy())
    assert res[0]["timing"]["launches"] < 30                  # for all 64 histories together


@pytest.mark.gpu
def test_long_histories_take_the_global_memory_paths():
    """Pools beyond 40 KB are gathered from global memory by the batched kernel and re-read by the calibration
    instead of being staged in shared memory: same results as one run per history / as the restatement."""
    from oracle import garch_fit
    from montecarlo_b200 import calibrate_garch, garch_monte_carlo_from_close
    lengths = np.array([9000, 6500, 700], dtype=np.uint32)
    close = np.full((3, 9000), 5.0)
    for b, n in enumerate(lengths):
        close[b, :n] = 30.0 * np.exp(np.cumsum(garch_fit.simulate_garch(int(n), 4e-6, 0.07, 0.9, seed=70 + b)))
    loop = garch_monte_carlo_from_close(close, 600, 15, lengths=lengths, seed=2, batched=False)
    batch = garch_monte_carlo_from_close(close, 600, 15, lengths=lengths, seed=2, batched=True)
    for a, b in zip(loop, batch):
        assert np.array_equal(a["paths"], b["paths"]) and a["percentiles"] == b["percentiles"]
    fits = calibrate_garch(close, lengths=lengths)
    for b, n in enumerate(lengths):
        exp = garch_fit.calibrate(port.history_inputs(close[b, :n])["returns"])
        assert abs(fits[b]["objective"] - exp["objective"]) <= 1e-9 * abs(exp["objective"])
        assert abs(fits[b]["alpha"] - exp["alpha"]) <= 2e-4 and abs(fits[b]["beta"] - exp["beta"]) <= 2e-4
