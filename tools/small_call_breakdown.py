"""Where the wall time of the app's default call (1000 x 20 through garch_monte_carlo) goes, step by step (median of 300)."""
import os, sys, time, statistics, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd, torch
from montecarlo_b200 import garch_monte_carlo, api, _lib as L
from montecarlo_b200.engine import get_engine, _CudaStageBackend
from montecarlo_b200.spec import SimSpec, make_params, make_portfolio, REF_TERMINAL_PCTS, REF_PATH_PCTS, REF_TAIL_PCT

rs = np.random.RandomState(0)
df = pd.DataFrame({"log_ret": rs.normal(0.0, 0.02, 1250), "rvol_20": 0.02 * np.sqrt(252)})
n, t = 1000, 20
eng = get_engine(0)
for _ in range(20):
    garch_monte_carlo(df, n, t, 100.0, seed=1)
T = {}
def lap(name, t0):
    T.setdefault(name, []).append((time.perf_counter() - t0) * 1e6)
for _ in range(300):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); returns, rvol = api._history_inputs(df); lap("history_inputs (pandas -> numpy)", t0)
    t0 = time.perf_counter(); pool = api.bootstrap_pool(returns); var0 = rvol ** 2 / 252; lap("bootstrap_pool", t0)
    t0 = time.perf_counter()
    spec = SimSpec(n_paths=n, n_steps=t, s0=100.0, var0=var0, dtype="float64", seed=1, store_paths=True, fan_chart=True,
                   pcts=REF_TERMINAL_PCTS, path_pcts=REF_PATH_PCTS, tail_pct=REF_TAIL_PCT)
    lap("SimSpec", t0)
    t0 = time.perf_counter(); pool_t = eng.to_device(pool, torch.float64); spec.pool_len = pool_t.numel(); lap("pool -> device", t0)
    t0 = time.perf_counter()
    with torch.cuda.device(eng.device):
        terminal = torch.empty(n, dtype=torch.float64, device=eng.device)
        paths = torch.empty((t + 1, n), dtype=torch.float64, device=eng.device)
        lap("torch.empty x2 (+ device ctx)", t0)
        t0 = time.perf_counter(); params = make_params(spec, 0, n, n_ranks=1); pf = make_portfolio(spec); lap("make_params", t0)
        t0 = time.perf_counter()
        res = L.McResult(); fan = np.zeros((5, t + 1)); sp = eng.stream_ptr()
        lap("result holders + stream_ptr", t0)
        t0 = time.perf_counter()
        rc = eng.lib.mc_run(eng.handle, C.byref(params), None, C.c_void_p(pool_t.data_ptr()), None, C.c_void_p(paths.data_ptr()),
                            C.c_void_p(terminal.data_ptr()), C.byref(res), None, C.c_void_p(fan.ctypes.data), sp)
        lap("mc_run (ctypes, incl. device wait)", t0)
    t0 = time.perf_counter(); out = eng._pack(spec, res, None, fan, terminal, paths, 0); lap("_pack", t0)
    t0 = time.perf_counter(); d = api._result_dict(out, 100.0, "numpy"); lap("_result_dict (incl. paths D2H 168 KB)", t0)
    t0 = time.perf_counter(); garch_monte_carlo(df, n, t, 100.0, seed=1); lap("== whole garch_monte_carlo call", t0)
tot = 0.0
for k, v in T.items():
    m = statistics.median(v)
    if not k.startswith("=="):
        tot += m
    print(f"{m:8.1f} us  {k}")
print(f"{tot:8.1f} us  sum of the steps")
