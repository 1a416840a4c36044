"""BASELINE config 1 (the reference's own shape: 10k x 252, fp64, full result dict) and the app's default
call (1000 x 20) through the drop-in function: wall time per call."""
import os, sys, time, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd, torch
from montecarlo_b200 import garch_monte_carlo
rs = np.random.RandomState(0)
df = pd.DataFrame({"log_ret": rs.normal(0.0, 0.02, 1250), "rvol_20": 0.02 * np.sqrt(252)})
for n, t, kw in ((1000, 20, {}), (5000, 60, {}), (10_000, 252, {}), (10_000, 252, dict(return_paths="device")),
                 (10_000, 252, dict(dtype="float32", return_paths="device")), (100_000, 252, dict(return_paths="device"))):
    for _ in range(3):
        garch_monte_carlo(df, n, t, 100.0, seed=1, **kw)
    ts = []
    for _ in range(20):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = garch_monte_carlo(df, n, t, 100.0, seed=1, **kw)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(f"{n:7d} x {t:3d} {kw}: median {statistics.median(ts):.3f} ms  min {min(ts):.3f} ms  "
          f"device total {r['timing']['total_ms']:.3f} ms sim {r['timing']['sim_ms']:.3f} ms launches {r['timing']['launches']}")
