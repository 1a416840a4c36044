"""Headless stand-in for the dashboard around the hot path -- TEST INFRASTRUCTURE.

The reference reaches `garch_monte_carlo` from exactly one place, Single-Stock mode (app.py:1234-1241), and
reads the result in exactly one place, the Monte Carlo tab (app.py:944-1000).  Streamlit, plotly and yfinance
are not installed here, so this module re-enacts both ends without them:

* `synthetic_ohlcv` / `hot_path_columns` build the frame the call site hands over: OHLCV rows, the two
  columns the routine reads (`log_ret` app.py:224, `rvol_20` app.py:233) and the row dropping of app.py:392-393
  (the 200-day moving average of app.py:252 leaves the first 199 rows NaN).
* `run_single_stock` makes the call the way app.py:1240 does (five positional arguments, the status callback).
* `render_monte_carlo_tab` evaluates every expression the tab evaluates on `mc_results` and returns what
  would have been drawn: the four metric strings, the percentile table and the plot traces.

Any implementation with the reference's signature can be passed in, so the same driver runs over the CUDA
drop-in (GPU tests) and over the unmodified reference / the NumPy port (CPU tests, which validate the shim).
"""
from __future__ import annotations

import json

import numpy as np
import pandas as pd

from oracle.ref_loader import synthetic_ohlcv  # noqa: F401  (re-exported for the tests)

SMA_WARMUP_ROWS = 199          # close.rolling(200).mean() is NaN on rows 0..198 (app.py:252), dropped at app.py:393
PERCENTILE_LABELS = ("1%", "5%", "10%", "25%", "50%", "75%", "90%", "95%", "99%")
TERMINAL_KEYS = ("p1", "p5", "p10", "p25", "p50", "p75", "p90", "p95", "p99")
BAND_COLOURS = {"p5": "red", "p25": "orange", "p50": "blue", "p75": "green", "p95": "darkgreen"}


def hot_path_columns(ohlcv: pd.DataFrame) -> pd.DataFrame:
    """The part of `add_all_features` the routine depends on, with the same pandas calls."""
    data = ohlcv.copy()
    close = data["Close"]
    data["log_ret"] = np.log(close / close.shift(1))                      # app.py:224
    data["rvol_20"] = data["log_ret"].rolling(20).std() * np.sqrt(252)     # app.py:233
    data["sma_200"] = close.rolling(200).mean()                            # app.py:252 -- sets the warm-up
    data = data.replace([np.inf, -np.inf], 0).dropna()                     # app.py:392-393
    return data.drop(columns=["sma_200"])


class _Recorder:
    """Collects what the tab would have shown."""

    def __init__(self):
        self.metrics, self.traces, self.table, self.layout = {}, [], None, None


def render_monte_carlo_tab(mc_results, current_price: float, predicted_pct_return: float = 1.5):
    """app.py:944-1000 on a recorder.  Returns None when the tab is skipped (falsy result, app.py:945)."""
    if not mc_results:
        return None
    out = _Recorder()
    out.metrics["Expected Return"] = f"{mc_results['expected_return']:.2f}%"      # app.py:950
    out.metrics["Prob. of Profit"] = f"{mc_results['prob_profit']:.1f}%"          # app.py:951
    out.metrics["VaR (95%)"] = f"{mc_results['var_95']:.2f}%"                     # app.py:952
    out.metrics["CVaR (95%)"] = f"{mc_results['cvar_95']:.2f}%"                   # app.py:953
    pct = mc_results["percentiles"]
    prices = [pct[k] for k in TERMINAL_KEYS]                                       # app.py:959-960
    table = pd.DataFrame({"Percentile": list(PERCENTILE_LABELS), "Price": prices,
                          "Return": [(p / current_price - 1) * 100 for p in prices]})   # app.py:961-963
    out.table = table
    out.table_text = [("${:.2f}".format(p), "{:+.2f}%".format(r)) for p, r in zip(table["Price"], table["Return"])]
    path_pct = mc_results["path_percentiles"]
    for key, colour in BAND_COLOURS.items():                                       # app.py:974-980
        y = path_pct[key]
        out.traces.append({"y": json.loads(json.dumps(np.asarray(y).tolist())), "mode": "lines",
                           "name": key.replace("p", "") + "th %ile",
                           "line": {"color": colour, "width": 2 if key == "p50" else 1}})
    ml_target = current_price * (1 + predicted_pct_return / 100)                   # app.py:983
    out.traces.append({"x": [len(path_pct["p50"]) - 1], "y": [ml_target], "mode": "markers", "name": "ML Target"})
    out.layout = {"title": "Price Projection with Percentile Bands", "xaxis_title": "Days", "yaxis_title": "Price ($)"}
    return out


def run_single_stock(garch_monte_carlo, ohlcv: pd.DataFrame, mc_sims: int = 1000, forecast_days: int = 20):
    """app.py:1207-1241 without the model training in between: returns (frame, price, messages, result, tab)."""
    df = hot_path_columns(ohlcv)
    current_price = df["Close"].iloc[-1]                                           # app.py:1213
    messages = []

    def update_status(msg):                                                        # app.py:1220-1221
        messages.append(msg)

    mc_results = garch_monte_carlo(df, mc_sims, forecast_days, current_price, update_status)   # app.py:1240
    return df, current_price, messages, mc_results, render_monte_carlo_tab(mc_results, current_price)
