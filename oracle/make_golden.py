"""Generate tests/golden/ref_*.npz by running the UNMODIFIED reference function
(/root/reference/app.py:586-657, loaded by oracle/ref_loader.py) in the build
container.  Re-run with:  python -m oracle.make_golden

The fixtures hold inputs and reference outputs only (no reference source).  Each
case also asserts, at generation time, that the oracle port (oracle/mc_numpy.py)
reproduces the reference bit-for-bit on the same shocks.
"""
from __future__ import annotations

import os

import numpy as np

from . import mc_numpy as port
from . import ref_loader

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                          "tests", "golden")
SCALARS = ("expected_return", "expected_vol", "prob_profit", "var_95", "cvar_95")


def _pack(res):
    out = {k: np.float64(res[k]) for k in SCALARS}
    out["percentiles"] = np.array([res["percentiles"][f"p{p}"] for p in port.REF_TERMINAL_PCTS])
    out["path_percentiles"] = np.stack([res["path_percentiles"][f"p{p}"] for p in port.REF_PATH_PCTS])
    return out


def _check_port(res, mine):
    assert np.array_equal(res["paths"], mine["paths"])
    for k in SCALARS:
        assert res[k] == mine[k], k
    for k, v in res["percentiles"].items():
        assert v == mine["percentiles"][k], k
    for k, v in res["path_percentiles"].items():
        assert np.array_equal(v, mine["path_percentiles"][k]), k


def injected_case(name, n_sims, n_days, s0, frame_seed, heavy, idx_seed, normal=False):
    df = ref_loader.synthetic_frame(1250 if not heavy else 777, frame_seed, heavy)
    returns = df["log_ret"].values
    rvol_last = float(df["rvol_20"].iloc[-1])
    rs = np.random.RandomState(idx_seed)
    kw = {}
    if normal:   # normal shocks through the reference's own GARCH loop
        kw["shocks"] = rs.standard_normal((n_days, n_sims))
    else:
        kw["indices"] = rs.randint(0, len(returns), size=(n_days, n_sims)).astype(np.int32)
    res = ref_loader.run_reference_injected(df, n_sims, n_days, s0, **kw)
    mine = port.garch_monte_carlo_port(returns, rvol_last, n_sims, n_days, s0, **kw)
    _check_port(res, mine)
    data = dict(returns=returns, rvol_last=rvol_last, s0=s0, n_sims=n_sims, n_days=n_days,
                paths=res["paths"], **kw, **_pack(res))
    np.savez_compressed(os.path.join(GOLDEN_DIR, name), **data)
    print(name, "ok", res["paths"].shape)


def seeded_case(name, n_sims, n_days, s0, np_seed):
    """Reference run under np.random.seed (BASELINE config 1).  Only outputs are
    stored; the test regenerates the draws with the same legacy RandomState."""
    df = ref_loader.synthetic_frame(1250, 0, False)
    returns = df["log_ret"].values
    rvol_last = float(df["rvol_20"].iloc[-1])
    fn = ref_loader.load_reference_function()
    np.random.seed(np_seed)
    res = fn(df, n_sims, n_days, s0)
    mine = port.garch_monte_carlo_port(returns, rvol_last, n_sims, n_days, s0,
                                       rng=np.random.RandomState(np_seed))
    _check_port(res, mine)
    data = dict(returns=returns, rvol_last=rvol_last, s0=s0, n_sims=n_sims, n_days=n_days,
                np_seed=np_seed, final_every_97=res["final_prices"][::97], **_pack(res))
    np.savez_compressed(os.path.join(GOLDEN_DIR, name), **data)
    print(name, "ok")


def inputs_case(name, cases):
    """the routine's inputs from raw prices.  Runs the UNMODIFIED `add_all_features`
    (app.py:210-395) on synthetic OHLCV histories of different lengths, takes from its frame what
    `garch_monte_carlo` reads (app.py:592, 606, 609) and the call site's current price (app.py:1213), and
    stores them next to the close prices.  The port (`mc_numpy.history_inputs`) must agree bit for bit."""
    import warnings
    features = ref_loader.load_reference_features()
    n_max = max(c[0] for c in cases)
    close = np.zeros((len(cases), n_max))
    lengths = np.zeros(len(cases), dtype=np.uint32)
    pool = np.zeros((len(cases), n_max))
    pool_len = np.zeros(len(cases), dtype=np.int64)
    pool_std, var0, s0 = (np.zeros(len(cases)) for _ in range(3))
    for i, (n_rows, seed, heavy) in enumerate(cases):
        ohlcv = ref_loader.synthetic_ohlcv(n_rows, seed, 40.0 + 35.0 * i, heavy)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")                      # the reference silences them too (app.py:15)
            frame = features(ohlcv)
        close[i, :n_rows], lengths[i] = ohlcv["Close"].values, n_rows
        mine = port.history_inputs(ohlcv["Close"].values)
        if len(frame) == 0:
            assert mine is None
            continue
        returns = frame["log_ret"].dropna().values                # app.py:592
        assert len(frame) == n_rows - port.SMA_WARMUP_ROWS and len(returns) == len(frame)
        std = np.std(returns)
        pool_len[i], pool_std[i] = len(returns), std
        pool[i, :len(returns)] = returns / (std + 1e-10)         # app.py:609
        var0[i] = (frame["rvol_20"].iloc[-1] ** 2) / 252          # app.py:606
        s0[i] = ohlcv["Close"].iloc[-1]                           # app.py:1213
        assert np.array_equal(mine["pool"], pool[i, :pool_len[i]]) and mine["var0"] == var0[i]
        assert mine["s0"] == s0[i] and mine["pool_std"] == pool_std[i]
    np.savez_compressed(os.path.join(GOLDEN_DIR, name), close=close, lengths=lengths, pool=pool, pool_len=pool_len,
                        pool_std=pool_std, var0=var0, s0=s0, first_row=port.SMA_WARMUP_ROWS)
    print(name, "ok", close.shape, pool_len)


def main():
    if not ref_loader.reference_available():
        raise SystemExit("reference not present: golden fixtures can only be made in the build container")
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    injected_case("ref_small_bootstrap.npz", 64, 20, 100.0, 0, False, 11)
    injected_case("ref_ragged_heavytail.npz", 257, 33, 42.5, 3, True, 12)
    injected_case("ref_normal_through_garch.npz", 128, 40, 100.0, 0, False, 13, normal=True)
    seeded_case("ref_c1_seed42.npz", 10_000, 252, 100.0, 42)
    inputs_case("ref_inputs_from_close.npz", [(1260, 11, False), (900, 12, True), (320, 13, False), (260, 14, False),
                                              (199, 15, False), (1259, 16, True)])
    # short-history guard (app.py:594-595)
    df = ref_loader.synthetic_frame(99, 0, False)
    assert ref_loader.load_reference_function()(df, 10, 5, 100.0) is None
    assert port.garch_monte_carlo_port(df["log_ret"].values, 0.3, 10, 5, 100.0) is None
    print("short-history guard ok")


if __name__ == "__main__":
    main()
