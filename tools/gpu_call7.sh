set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -12 gpurun_out/pytest.log
python tools/c1_latency.py 2>&1 | tail -20
python - <<'PY'
import cProfile, pstats, numpy as np, pandas as pd, torch
from montecarlo_b200 import garch_monte_carlo
rs = np.random.RandomState(0)
df = pd.DataFrame({"log_ret": rs.normal(0.0, 0.02, 1250), "rvol_20": 0.02 * np.sqrt(252)})
for _ in range(5): garch_monte_carlo(df, 1000, 20, 100.0, seed=1)
pr = cProfile.Profile(); pr.enable()
for _ in range(200): garch_monte_carlo(df, 1000, 20, 100.0, seed=1)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
PY
