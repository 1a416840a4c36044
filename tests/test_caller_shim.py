"""the dashboard's call site and Monte Carlo tab (app.py:1234-1241, 944-1000) run headless.

CPU tests validate the shim itself against the unmodified reference (when /root/reference is present) and the
NumPy port; GPU tests run the same driver over the CUDA drop-in."""
import numpy as np
import pytest

import caller_shim as shim
from oracle import mc_numpy as port
from oracle import ref_loader

STATUS = ["Running Monte Carlo simulation..."]          # app.py:597-598


def _port_as_drop_in(df, n_sims, n_days, current_price, progress_callback=None):
    returns = df["log_ret"].dropna().values               # app.py:592
    if len(returns) < 100:
        return None
    if progress_callback:
        progress_callback(STATUS[0])
    return port.garch_monte_carlo_port(returns, df["rvol_20"].iloc[-1], n_sims, n_days, current_price)


def _check_tab(tab, price, forecast_days):
    assert set(tab.metrics) == {"Expected Return", "Prob. of Profit", "VaR (95%)", "CVaR (95%)"}
    for text in tab.metrics.values():
        float(text.rstrip("%"))
    assert list(tab.table["Percentile"]) == list(shim.PERCENTILE_LABELS)
    assert tab.table["Price"].is_monotonic_increasing and np.all(np.isfinite(tab.table["Price"]))
    assert all(p.startswith("$") and r.endswith("%") for p, r in tab.table_text)
    bands = tab.traces[:5]
    assert [t["name"] for t in bands] == ["5th %ile", "25th %ile", "50th %ile", "75th %ile", "95th %ile"]
    for t in bands:
        assert len(t["y"]) == forecast_days + 1 and t["y"][0] == pytest.approx(price, rel=1e-6)
    stacked = np.array([t["y"] for t in bands])
    assert np.all(np.diff(stacked, axis=0) >= 0)          # the bands never cross
    assert tab.traces[5]["x"] == [forecast_days]          # app.py:986


def test_frame_has_what_the_routine_reads():
    ohlcv = shim.synthetic_ohlcv()
    df = shim.hot_path_columns(ohlcv)
    assert len(df) == len(ohlcv) - shim.SMA_WARMUP_ROWS
    assert {"log_ret", "rvol_20", "Close"} <= set(df.columns) and not df.isna().any().any()


def test_shim_over_the_numpy_port():
    np.random.seed(3)
    df, price, messages, res, tab = shim.run_single_stock(_port_as_drop_in, shim.synthetic_ohlcv(), 1000, 20)
    assert messages == STATUS
    _check_tab(tab, price, 20)


@pytest.mark.skipif(not ref_loader.reference_available(), reason="/root/reference exists only in the build container")
def test_shim_over_the_unmodified_reference_equals_the_port():
    ref = ref_loader.load_reference_function()
    ohlcv = shim.synthetic_ohlcv(heavy_tail=True)
    np.random.seed(9)
    _, price, messages, _, tab_ref = shim.run_single_stock(ref, ohlcv, 500, 5)
    np.random.seed(9)
    _, _, _, _, tab_port = shim.run_single_stock(_port_as_drop_in, ohlcv, 500, 5)
    assert messages == STATUS
    _check_tab(tab_ref, price, 5)
    assert tab_ref.metrics == tab_port.metrics and tab_ref.table_text == tab_port.table_text
    assert tab_ref.traces == tab_port.traces


def test_short_history_skips_the_tab():
    _, _, messages, res, tab = shim.run_single_stock(_port_as_drop_in, shim.synthetic_ohlcv(n_rows=260), 1000, 20)
    assert res is None and tab is None and messages == []


# ------------------------------------------------------------------------------------------------ CUDA drop-in
@pytest.mark.gpu
@pytest.mark.parametrize("mc_sims,forecast_days", [(500, 5), (1000, 20), (5000, 60)])   # slider ranges app.py:1189-1195
def test_dashboard_over_the_cuda_drop_in(mc_sims, forecast_days):
    from montecarlo_b200 import garch_monte_carlo
    ohlcv = shim.synthetic_ohlcv(seed=mc_sims)
    df, price, messages, res, tab = shim.run_single_stock(garch_monte_carlo, ohlcv, mc_sims, forecast_days)
    assert messages == STATUS
    _check_tab(tab, price, forecast_days)
    # the numbers on screen are the reference's own reductions of the returned paths (app.py:636-656)
    exp = port.summarize(np.asarray(res["paths"]), price)
    again = shim.render_monte_carlo_tab(exp, price)
    assert tab.metrics == again.metrics and tab.table_text == again.table_text and tab.traces == again.traces


@pytest.mark.gpu
def test_dashboard_numbers_agree_with_the_reference_model():
    """Same frame through the port (MT19937 draws) and the drop-in (Philox draws): different streams, same law."""
    from montecarlo_b200 import garch_monte_carlo
    ohlcv = shim.synthetic_ohlcv(heavy_tail=True)
    np.random.seed(21)
    _, price, _, ref, _ = shim.run_single_stock(_port_as_drop_in, ohlcv, 5000, 20)
    _, _, _, res, _ = shim.run_single_stock(lambda *a: garch_monte_carlo(*a, seed=21), ohlcv, 5000, 20)
    sd = np.std(ref["final_prices"])
    n = 5000
    assert abs(res["percentiles"]["p50"] - ref["percentiles"]["p50"]) < 4.5 * 1.2533 * sd * np.sqrt(2.0 / n)
    assert abs(res["expected_return"] - ref["expected_return"]) < 4.5 * sd / price * 100 * np.sqrt(2.0 / n)
    assert abs(res["prob_profit"] - ref["prob_profit"]) < 4.5 * 50 * np.sqrt(2.0 / n)
    assert abs(res["var_95"] - ref["var_95"]) < 0.15 * abs(ref["var_95"]) + 0.2


@pytest.mark.gpu
def test_short_history_skips_the_tab_on_the_drop_in():
    from montecarlo_b200 import garch_monte_carlo
    _, _, messages, res, tab = shim.run_single_stock(garch_monte_carlo, shim.synthetic_ohlcv(n_rows=260), 1000, 20)
    assert res is None and tab is None and messages == []
