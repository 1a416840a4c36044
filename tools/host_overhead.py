"""Where does a C2 job's time go?  device sim / device total (events inside mc_run) / wall of simulate()."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from montecarlo_b200.engine import get_engine
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
eng = get_engine(0)
spec, pool = bench.make_spec(wl, 1)
pool_dev = eng.to_device(pool, torch.float32) if pool is not None else None
for _ in range(5):
    eng.simulate(spec, pool=pool_dev)
torch.cuda.synchronize()
rows = []
for _ in range(20):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = eng.simulate(spec, pool=pool_dev)
    t1 = time.perf_counter()
    rows.append((out.sim_ms, out.total_ms, (t1 - t0) * 1e3))
import statistics
for i, name in enumerate(("sim_ms", "device_total_ms", "wall_simulate_ms")):
    print(name, round(statistics.median(r[i] for r in rows), 4))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(20):
    eng.simulate(spec, pool=pool_dev)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
