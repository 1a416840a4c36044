"""Load the UNMODIFIED reference function for pinning the oracle (TEST INFRASTRUCTURE).

`import app` is impossible (streamlit/yfinance/xgboost/plotly are absent and
app.py:17 runs `st.set_page_config` at import), so the FunctionDef named
`garch_monte_carlo` (app.py:586-657) is taken out of the parsed module and
executed in a namespace holding only what it uses: np, pd, Optional, Dict.

/root/reference exists only in the build container: nothing that runs on the GPU
box may call this.  `make_golden.py` uses it to write tests/golden/ref_*.npz.
"""
from __future__ import annotations

import ast
import os
from typing import Dict, Optional

import numpy as np

DEFAULT_REFERENCE = os.environ.get("MC_REFERENCE", "/root/reference/app.py")


# the function as it travels to the GPU box: extracted from the reference by `write_extracted_copy` (called from
# __graft_entry__.build in the build container) into a git-ignored directory; never committed
EXTRACTED_COPY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "garch_monte_carlo_from_app_py.py")


def reference_available(path: str = DEFAULT_REFERENCE) -> bool:
    return os.path.isfile(path)


def reference_path():
    """The reference itself if it is mounted here, else the extracted copy, else None."""
    if os.path.isfile(DEFAULT_REFERENCE):
        return DEFAULT_REFERENCE
    return EXTRACTED_COPY if os.path.isfile(EXTRACTED_COPY) else None


def write_extracted_copy(path: str = DEFAULT_REFERENCE) -> str:
    """oracle/_ref/: the FunctionDef `garch_monte_carlo` (app.py:586-657) exactly as the reference has it, unparsed
    from its AST, so that the GPU box (which has no /root/reference) can time the UNMODIFIED routine on its host."""
    with open(path, "r") as fh:
        src = fh.read()
    tree = ast.parse(src, filename=path)
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "garch_monte_carlo")
    text = ast.get_source_segment(src, node)
    os.makedirs(os.path.dirname(EXTRACTED_COPY), exist_ok=True)
    with open(EXTRACTED_COPY, "w") as fh:
        fh.write("# extracted verbatim from the reference's app.py:%d-%d by oracle/ref_loader.write_extracted_copy -- not "
                 "part of this repository's sources (git-ignored)\n" % (node.lineno, node.end_lineno))
        fh.write(text + "\n")
    return EXTRACTED_COPY


def load_reference_function(path: str = DEFAULT_REFERENCE, np_module=np):
    import pandas as pd
    with open(path, "r") as fh:
        tree = ast.parse(fh.read(), filename=path)
    node = next(n for n in tree.body
                if isinstance(n, ast.FunctionDef) and n.name == "garch_monte_carlo")
    module = ast.Module(body=[node], type_ignores=[])
    ns = {"np": np_module, "pd": pd, "Optional": Optional, "Dict": Dict}
    exec(compile(module, path, "exec"), ns)
    return ns["garch_monte_carlo"]


def load_reference_features(path: str = DEFAULT_REFERENCE):
    """The unmodified `add_all_features` (app.py:210-395) with the indicator helpers it calls
    (`compute_rsi`, `compute_atr`, `compute_macd`, app.py:180-208): the producer of the two columns the
    routine reads, `log_ret` (app.py:224) and `rvol_20` (app.py:233), and of the row dropping (app.py:392-393)."""
    import typing

    import pandas as pd
    with open(path, "r") as fh:
        tree = ast.parse(fh.read(), filename=path)
    nodes = [n for n in tree.body if isinstance(n, ast.FunctionDef)
             and (n.name.startswith("compute_") or n.name == "add_all_features")]
    ns = {"np": np, "pd": pd, "Optional": Optional, "Dict": Dict, "Tuple": typing.Tuple, "List": typing.List}
    exec(compile(ast.Module(body=nodes, type_ignores=[]), path, "exec"), ns)
    return ns["add_all_features"]


class _ChoiceProxy:
    """Stands in for `np.random` inside the reference so that a supplied matrix of
    shocks (or pool indices) is what `np.random.choice` (app.py:620) returns."""

    def __init__(self, shocks=None, indices=None):
        self._shocks, self._indices, self._t = shocks, indices, 0

    def choice(self, pool, size=None):
        t, self._t = self._t, self._t + 1
        if self._shocks is not None:
            return np.asarray(self._shocks[t], dtype=np.float64)
        return np.asarray(pool)[np.asarray(self._indices[t])]


class _NumpyProxy:
    def __init__(self, random):
        self.random = random

    def __getattr__(self, name):
        return getattr(np, name)


def run_reference_injected(df, n_sims, n_days, s0, *, shocks=None, indices=None,
                           path: str = DEFAULT_REFERENCE):
    """Run the unmodified reference with its RNG replaced by a supplied matrix."""
    fn = load_reference_function(path, _NumpyProxy(_ChoiceProxy(shocks, indices)))
    return fn(df, n_sims, n_days, s0)


def synthetic_frame(n_rows: int = 1250, seed: int = 0, heavy_tail: bool = False):
    """SURVEY 8c synthetic input: log_ret ~ N(0, 0.02^2) (or scaled Student-t4),
    rvol_20 = 0.02*sqrt(252)."""
    import pandas as pd
    rs = np.random.RandomState(seed)
    if heavy_tail:
        r = rs.standard_t(4, n_rows) * 0.02 / np.sqrt(2.0)
    else:
        r = rs.normal(0.0, 0.02, n_rows)
    return pd.DataFrame({"log_ret": r, "rvol_20": np.full(n_rows, 0.02 * np.sqrt(252))})


def synthetic_ohlcv(n_rows: int = 1260, seed: int = 11, s0: float = 120.0, heavy_tail: bool = False):
    """Five years of business days, the shape `get_stock_data(ticker, period="5y")` hands over (app.py:1207)."""
    import pandas as pd
    rs = np.random.RandomState(seed)
    r = rs.standard_t(4, n_rows) * 0.012 if heavy_tail else rs.normal(3e-4, 0.017, n_rows)
    close = s0 * np.exp(np.cumsum(r))
    spread = np.abs(rs.normal(0.0, 0.006, n_rows))
    idx = pd.bdate_range("2020-01-01", periods=n_rows)
    return pd.DataFrame({"Open": close * (1 - 0.3 * spread), "High": close * (1 + spread), "Low": close * (1 - spread),
                         "Close": close, "Volume": rs.randint(1_000_000, 9_000_000, n_rows).astype(np.float64)}, index=idx)
