"""Device timeline of one job (kernel start / duration / gap to the previous kernel) through torch.profiler (CUPTI):
what the closing stage of a job is made of once launch gaps are counted.   python tools/timeline.py [c2|c5|ref|c3]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from montecarlo_b200.engine import get_engine

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
eng = get_engine(0)
if os.environ.get("MC_SIDE_CAP"):          # experiments: capacity (values) of a warp's side list
    eng.debug_set_tuning(side_cap=int(os.environ["MC_SIDE_CAP"]))
spec, pool = bench.make_spec(wl, 1)
pool_dev = eng.to_device(pool, torch.float32) if pool is not None else None
flush = torch.empty(256 << 20, dtype=torch.uint8, device=eng.device)
for _ in range(5):
    out = eng.simulate(spec, pool=pool_dev)
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        flush.zero_()
        torch.cuda.synchronize()
        out = eng.simulate(spec, pool=pool_dev)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
# last job: from the last simulation kernel on
sims = [i for i, e in enumerate(evs) if "sim32_kernel" in e.name or "portfolio_kernel" in e.name or "sim_kernel" in e.name]
start = sims[-1]
prev_end = None
t0 = evs[start].time_range.start
tot_k = tot_gap = 0.0
for e in evs[start:]:
    s, d = e.time_range.start, e.time_range.end - e.time_range.start
    gap = (s - prev_end) if prev_end is not None else 0.0
    print(f"{(s - t0):9.1f} us  dur {d:8.1f}  gap {gap:6.1f}  {e.name[:90]}")
    tot_k += d; tot_gap += max(gap, 0.0)
    prev_end = e.time_range.end
print(f"kernels {tot_k:.1f} us, gaps {tot_gap:.1f} us, span {(prev_end - t0):.1f} us")
