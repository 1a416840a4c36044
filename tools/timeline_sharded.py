"""Device timeline of one SHARDED job on rank 0 (torch.profiler / CUPTI): what the two peer-memory exchanges cost.
   python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/timeline_sharded.py [c2|c4|ref]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from torch.profiler import profile, ProfilerActivity
import bench
from montecarlo_b200.engine import get_engine

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
group = dist.group.WORLD
eng = get_engine(local)
spec, pool = bench.make_spec(wl, world)
pool_dev = eng.to_device(pool, torch.float32) if pool is not None else None
flush = torch.empty(256 << 20, dtype=torch.uint8, device=eng.device)
for _ in range(5):
    out = eng.simulate(spec, pool=pool_dev, group=group)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(4):
        flush.zero_()
        torch.cuda.synchronize()
        dist.barrier(group)
        torch.cuda.synchronize()
        out = eng.simulate(spec, pool=pool_dev, group=group)
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    sims = [i for i, e in enumerate(evs) if "sim32_kernel" in e.name or "portfolio_kernel" in e.name or "sim_kernel" in e.name]
    for start in sims[-3:]:
        t0 = evs[start].time_range.start
        prev_end = None
        for e in evs[start:start + 4]:
            s, d = e.time_range.start, e.time_range.end - e.time_range.start
            gap = (s - prev_end) if prev_end is not None else 0.0
            print(f"{(s - t0):9.1f} us  dur {d:8.1f}  gap {gap:6.1f}  {e.name[:70]}")
            prev_end = e.time_range.end
        print(f"   span {(prev_end - t0):.1f} us   peers attached: {getattr(eng, '_peer_state', None)}")
dist.barrier()
dist.destroy_process_group()
