"""Host-side cost of the app's default call (1000 x 20 through garch_monte_carlo): the ctypes call mc_run itself,
Engine.simulate around it, and the drop-in function around that.  python tools/small_host_profile.py"""
import os, sys, time, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd, torch
from montecarlo_b200 import garch_monte_carlo
from montecarlo_b200.engine import get_engine

eng = get_engine(0)
lib = eng.lib
times = {"mc_run": []}
real = lib.mc_run


class Timed:
    def __call__(self, *a):
        t0 = time.perf_counter()
        rc = real(*a)
        times["mc_run"].append((time.perf_counter() - t0) * 1e6)
        return rc


class LibProxy:
    def __init__(self, lib):
        self._lib = lib
        self.mc_run = Timed()

    def __getattr__(self, k):
        return getattr(self._lib, k)


eng.lib = LibProxy(lib)
rs = np.random.RandomState(0)
df = pd.DataFrame({"log_ret": rs.normal(0.0, 0.02, 1250), "rvol_20": 0.02 * np.sqrt(252)})
for n, t in ((1000, 20), (10_000, 252)):
    for _ in range(10):
        garch_monte_carlo(df, n, t, 100.0, seed=1)
    times["mc_run"].clear()
    wall = []
    for _ in range(100):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = garch_monte_carlo(df, n, t, 100.0, seed=1)
        wall.append((time.perf_counter() - t0) * 1e6)
    print(f"{n} x {t}: garch_monte_carlo median {statistics.median(wall):.0f} us; mc_run (ctypes call incl. device wait) median "
          f"{statistics.median(times['mc_run']):.0f} us; device {r['timing']['total_ms'] * 1e3:.0f} us")
