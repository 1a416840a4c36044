"""Host-side logic and the C-ABI surface, without a GPU: the library loads, exports every
symbol include/mc_b200.h declares, and the ctypes structs match the C layout."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mc_b200.h")


@pytest.fixture(scope="module")
def lib():
    from montecarlo_b200.build import build_library
    from montecarlo_b200 import _lib
    build_library()
    return _lib.load_library()


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mc_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(lib):
    from montecarlo_b200 import _lib
    names = declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in mc_b200.h but not exported"
    assert sorted(_lib.EXPORTS) == names
    assert lib.mc_abi_version() == _lib.MC_ABI_VERSION


def test_struct_layout_matches_c(tmp_path):
    from montecarlo_b200 import _lib
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "mc_b200.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(mc_params), sizeof(mc_result),'
                   'sizeof(mc_portfolio), offsetof(mc_params, pcts), offsetof(mc_params, tail_pct),'
                   'offsetof(mc_result, tail_count), sizeof(mc_series_inputs), offsetof(mc_series_inputs, pool_len),'
                   'sizeof(mc_batch_params), offsetof(mc_batch_params, tail_pct));'
                   'return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    got = [C.sizeof(_lib.McParams), C.sizeof(_lib.McResult), C.sizeof(_lib.McPortfolio),
           _lib.McParams.pcts.offset, _lib.McParams.tail_pct.offset, _lib.McResult.tail_count.offset,
           C.sizeof(_lib.McSeriesInputs), _lib.McSeriesInputs.pool_len.offset,
           C.sizeof(_lib.McBatchParams), _lib.McBatchParams.tail_pct.offset]
    assert [int(v) for v in out] == got
    import numpy as np
    assert np.dtype(_lib.SERIES_INPUTS_DTYPE).itemsize == C.sizeof(_lib.McSeriesInputs)


def test_no_gpu_fails_loudly(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    assert lib.mc_create(0, C.byref(h)) != 0
    assert lib.mc_last_error(None)
    from montecarlo_b200.engine import Engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Engine(0)
    import pandas as pd
    from montecarlo_b200 import garch_monte_carlo
    df = pd.DataFrame({"log_ret": np.random.RandomState(0).normal(0, 0.02, 300), "rvol_20": 0.3})
    with pytest.raises(RuntimeError):
        garch_monte_carlo(df, 100, 5, 100.0)


def test_missing_library_raises(tmp_path, monkeypatch):
    code = ("import os, sys; sys.path.insert(0, %r); os.environ['MC_B200_LIB'] = %r\n"
            "from montecarlo_b200 import _lib\n"
            "try:\n    _lib.load_library()\nexcept RuntimeError as e:\n    print('RAISED', 'no CPU fallback' in str(e).replace('There is no CPU fallback', 'no CPU fallback'))\n"
            % (ROOT, str(tmp_path / "nope.so")))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, check=True).stdout
    assert "RAISED True" in out


def test_short_history_returns_none_without_gpu():
    import pandas as pd
    from montecarlo_b200 import garch_monte_carlo
    calls = []
    df = pd.DataFrame({"log_ret": np.r_[np.nan, np.zeros(99)], "rvol_20": 0.3})
    assert garch_monte_carlo(df, 100, 5, 100.0, calls.append) is None      # app.py:594-595
    assert calls == []                                                     # callback comes after the guard


def test_shard_range_partitions_exactly():
    from montecarlo_b200 import shard_range
    for n in (1, 7, 10_000_000, 100_000_001):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0
            assert sum(c for _, c in parts) == n
            for (o1, c1), (o2, _) in zip(parts, parts[1:]):
                assert o1 + c1 == o2
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1


def test_make_params_defaults_are_the_reference_constants():
    from montecarlo_b200 import SimSpec, make_params
    spec = SimSpec(n_paths=1000, n_steps=20, s0=100.0, var0=0.0004, pool_len=1250)
    p = make_params(spec)
    assert (p.omega, p.alpha, p.beta) == (0.00001, 0.10, 0.85)           # app.py:601-603
    assert list(p.pcts[:p.n_pcts]) == [1, 5, 10, 25, 50, 75, 90, 95, 99]  # app.py:639
    assert p.tail_pct == 5.0 and p.n_path_pcts == 0 and p.flags == 0
    assert p.struct_size == C.sizeof(type(p)) and p.n_paths == p.n_paths_global == 1000
    g = SimSpec.gbm(10, 4, 100.0, 0.05, 0.2, 1 / 252)
    pg = make_params(g, path_offset=5, n_local=3)
    assert pg.var_model == 1 and pg.shock == 1 and pg.n_paths == 3 and pg.path_offset == 5
    assert abs(pg.var0 - 0.04 / 252) < 1e-18 and abs(pg.drift - (0.05 - 0.02) / 252) < 1e-18
    with pytest.raises(ValueError):
        make_params(SimSpec(n_paths=10, n_steps=4, s0=1.0, fan_chart=True))
    with pytest.raises(ValueError):
        make_params(SimSpec(n_paths=10, n_steps=4, s0=1.0, pcts=(50, 5)))


def test_portfolio_spec_arrays():
    from montecarlo_b200 import PortfolioSpec, SimSpec, make_params
    from montecarlo_b200.spec import make_portfolio
    d = 16
    corr = np.full((d, d), 0.3) + 0.7 * np.eye(d)
    pf = PortfolioSpec(s0=np.full(d, 100.0), mu=0.05, sigma=0.15 + 0.01 * np.arange(d), weights=np.full(d, 1 / d),
                       corr=corr)
    spec = SimSpec(n_paths=100, n_steps=5, s0=0.0, shock="normal", var_model="const", portfolio=pf)
    p = make_params(spec)
    c = make_portfolio(spec)
    assert p.n_assets == d and abs(p.s0 - 100.0) < 1e-12
    chol = np.ctypeslib.as_array(c.h_chol, (d * d,)).reshape(d, d)
    assert np.allclose(chol @ chol.T, corr)


def test_fast_host_paths_equal_the_reference_expressions():
    """The drop-in's host side avoids pandas Series construction and np.std's dispatch overhead (they cost more than
    the simulation of the app's default call): same numbers as `df['log_ret'].dropna().values`, `df['rvol_20'].iloc[-1]`
    and `returns / (np.std(returns) + 1e-10)` (app.py:592, 606, 609), bit for bit."""
    import pandas as pd
    from montecarlo_b200 import api
    rs = np.random.RandomState(3)
    for n in (1, 2, 100, 1250, 4097, 100_003):
        r = rs.standard_t(4, n) * 0.013
        assert api._population_std(r) == float(np.std(r)), n
        assert np.array_equal(api.bootstrap_pool(r), r / (np.std(r) + 1e-10))
    assert np.array_equal(api.bootstrap_pool(np.arange(5, dtype=np.float32)), np.arange(5, dtype=np.float32) / (np.std(np.arange(5, dtype=np.float32)) + 1e-10))
    x = rs.normal(0, 0.02, 500)
    x[[0, 17, 499]] = np.nan
    for df in (pd.DataFrame({"log_ret": x, "rvol_20": np.linspace(0.1, 0.4, 500), "other": "a"}),
               pd.DataFrame({"rvol_20": np.linspace(0.1, 0.4, 500).astype(np.float32), "log_ret": x}),
               {"log_ret": pd.Series(x), "rvol_20": pd.Series(np.linspace(0.1, 0.4, 500))}):
        returns, rvol_last = api._history_inputs(df)
        col = df["log_ret"]
        assert np.array_equal(returns, col.dropna().values)
        assert rvol_last == float(df["rvol_20"].iloc[-1])


def test_fused_request_mirrors_what_mc_run_sharded_accepts():
    """Engine.fused_request decides, identically on every rank and without a GPU, whether a sharded run goes through
    `mc_run_sharded` (include/mc_b200.h: no fan chart, an explicit histogram range of <= 16384 bins, <= 16 ranks) or
    through the staged protocol."""
    from montecarlo_b200.engine import Engine
    from montecarlo_b200.spec import SimSpec
    gbm = lambda **kw: SimSpec.gbm(1000, 10, 100.0, 0.05, 0.2, 1 / 252, **kw)
    assert Engine.fused_request(gbm(), 8)
    assert Engine.fused_request(gbm(hist_bins=2048, hist_range=(50.0, 200.0)), 2)
    assert Engine.fused_request(gbm(hist_bins=16384, hist_range=(50.0, 200.0)), 16)
    assert not Engine.fused_request(gbm(hist_bins=2048), 2)                               # range from the data
    assert not Engine.fused_request(gbm(hist_bins=16385, hist_range=(50.0, 200.0)), 2)
    assert not Engine.fused_request(gbm(store_paths=True, fan_chart=True), 2)
    assert not Engine.fused_request(gbm(), 17)


def test_make_params_cache_hands_out_independent_copies():
    """The parameter struct of a job is cached by everything it is built from; callers get copies (the engine edits
    `flags` for its staged fall-back), invalid specs never get in, portfolio specs are always rebuilt."""
    import numpy as np
    from montecarlo_b200 import PortfolioSpec
    from montecarlo_b200.spec import SimSpec, make_params, _PARAMS_CACHE
    mk = lambda **kw: SimSpec.gbm(1234, 7, 100.0, 0.05, 0.2, 1 / 252, seed=5, hist_bins=16, hist_range=(50.0, 200.0), **kw)
    a = make_params(mk(), n_local=100, path_offset=7, n_ranks=3)
    b = make_params(mk(), n_local=100, path_offset=7, n_ranks=3)            # another spec object, same job: the cached struct
    assert bytes(a) == bytes(b) and a is not b
    a.flags |= 16
    assert make_params(mk(), n_local=100, path_offset=7, n_ranks=3).flags == b.flags
    c = make_params(mk(), n_local=101, path_offset=7, n_ranks=3)
    assert c.n_paths == 101 and b.n_paths == 100
    assert make_params(mk(dtype="float64")).dtype != b.dtype
    n_before = len(_PARAMS_CACHE)
    with pytest.raises(ValueError):
        make_params(SimSpec.gbm(10, 5, 100.0, 0.05, 0.2, pcts=(50.0, 5.0)))
    assert len(_PARAMS_CACHE) == n_before
    d = 2
    pf = PortfolioSpec(s0=np.full(d, 100.0), mu=0.05, sigma=np.full(d, 0.2), weights=np.full(d, 0.5), corr=np.eye(d), dt=1 / 252)
    sp = SimSpec(n_paths=10, n_steps=5, s0=100.0, shock="normal", var_model="const", portfolio=pf)
    assert make_params(sp).n_assets == d and len(_PARAMS_CACHE) == n_before
