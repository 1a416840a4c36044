"""Median duration of every kernel of a job over N jobs (torch.profiler / CUPTI): single timelines of the selection
kernels differ by +-10 % from job to job and box to box.   python tools/kernel_medians.py c5 [jobs]"""
import os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from montecarlo_b200.engine import get_engine

wl = sys.argv[1] if len(sys.argv) > 1 else "c5"
jobs = int(sys.argv[2]) if len(sys.argv) > 2 else 9
eng = get_engine(0)
spec, pool = bench.make_spec(wl, 1)
pool_dev = eng.to_device(pool, torch.float32) if pool is not None else None
flush = torch.empty(256 << 20, dtype=torch.uint8, device=eng.device)
for _ in range(4):
    out = eng.simulate(spec, pool=pool_dev)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(jobs):
        flush.zero_()
        torch.cuda.synchronize()
        out = eng.simulate(spec, pool=pool_dev)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
per_job, cur = [], None
for e in evs:
    if any(k in e.name for k in ("sim32_kernel", "portfolio_kernel", "sim_kernel<")):
        cur = []
        per_job.append(cur)
    if cur is not None and "elementwise" not in e.name and "fill" not in e.name.lower():
        cur.append(e)
per_job = [j for j in per_job if len(j) == len(per_job[-1])]
spans = [j[-1].time_range.end - j[0].time_range.start for j in per_job]
print(f"{wl}: {len(per_job)} jobs, span median {statistics.median(spans):.1f} us (min {min(spans):.1f}, max {max(spans):.1f})")
for k in range(len(per_job[-1])):
    d = [j[k].time_range.end - j[k].time_range.start for j in per_job]
    if statistics.median(d) >= 10.0:
        print(f"  {statistics.median(d):8.1f} us (min {min(d):7.1f} max {max(d):7.1f})  {per_job[-1][k].name[:80]}")
