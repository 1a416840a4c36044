"""Real-NCCL check of the sharding contract: the same GLOBAL job run on W GPUs (torchrun) must give,
on every rank, the very numbers one GPU gives for it -- exact order statistics, histogram, counts bit for bit,
moments to rounding.  Rank 0 also runs the job alone and compares.
   python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/nccl_parity.py"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from montecarlo_b200 import SimSpec
from montecarlo_b200.engine import get_engine
from oracle import mc_numpy as port

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = get_engine(local)
pool = port.bootstrap_pool(np.random.RandomState(5).standard_t(4, 1250) * 0.014)
cases = {
    "gbm 8,000,003 x 252 fp32 + histogram": (SimSpec.gbm(8_000_003, 252, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=11,
                                                        hist_bins=2048, hist_range=(20.0, 400.0)), None),
    "garch 3,000,001 x 60 fp32, stored paths + fan chart": (SimSpec(n_paths=3_000_001, n_steps=60, s0=100.0, var0=1.6e-4, dtype="float32",
                                                                  seed=12, store_paths=True, fan_chart=True), pool),
    "garch 400,003 x 33 fp64, stored paths + fan chart": (SimSpec(n_paths=400_003, n_steps=33, s0=42.5, var0=1.6e-4, dtype="float64",
                                                                seed=13, store_paths=True, fan_chart=True, hist_bins=64), pool),
    "garch 2,000,000 x 40 fp32, fan chart by regeneration": (SimSpec(n_paths=2_000_000, n_steps=40, s0=100.0, var0=1.6e-4, dtype="float32",
                                                                    seed=14, store_paths=True, fan_chart=True, fan_chunk_paths=100_000), pool),
}
from montecarlo_b200 import PortfolioSpec
d = 16
pf = PortfolioSpec(s0=[100.0] * d, mu=[0.05] * d, sigma=[0.15 + 0.01 * i for i in range(d)], weights=[1.0 / d] * d,
                   corr=0.3 * np.ones((d, d)) + 0.7 * np.eye(d))
cases["gbm 2,000,001 x 100 fp32, stored paths + fan chart"] = (
    SimSpec.gbm(2_000_001, 100, 100.0, 0.05, 0.2, 1 / 252, dtype="float32", seed=15, store_paths=True, fan_chart=True), None)
cases["16-asset portfolio 1,000,003 x 252 fp32"] = (
    SimSpec(n_paths=1_000_003, n_steps=252, s0=100.0, shock="normal", var_model="const", dtype="float32", seed=16, portfolio=pf), None)
cases["garch 10,000,019 x 252 fp32 (the reference's model at config-2 size)"] = (
    SimSpec(n_paths=10_000_019, n_steps=252, s0=100.0, var0=1.6e-4, dtype="float32", seed=17, hist_bins=512), pool)
report = {}
for name, (spec, p) in cases.items():
    shard = eng.simulate(spec, pool=p, group=dist.group.WORLD)
    mine = dict(pct=shard.pct, n=shard.n, n_above=shard.n_above, tail_count=shard.tail_count, mean=shard.mean, std=shard.std,
                tail_sum=shard.tail_sum, hist=None if shard.hist is None else shard.hist.tolist(),
                fan=None if shard.fan is None else {k: v.tolist() for k, v in shard.fan.items()})
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)
    if rank == 0:
        same_on_all_ranks = all(e == everyone[0] for e in everyone)
        import copy
        solo_spec = copy.copy(spec)
        solo_spec.fan_chunk_paths = 0 if spec.n_paths * (spec.n_steps + 1) < 4e8 else spec.fan_chunk_paths
        solo = eng.simulate(solo_spec, pool=p)
        exact = (solo.pct == shard.pct and solo.n == shard.n and solo.n_above == shard.n_above and solo.tail_count == shard.tail_count
                 and (solo.hist is None or np.array_equal(solo.hist, shard.hist))
                 and (solo.fan is None or all(np.array_equal(solo.fan[k], shard.fan[k]) for k in solo.fan)))
        close = (abs(solo.mean - shard.mean) <= 1e-12 * abs(solo.mean) and abs(solo.std - shard.std) <= 1e-10 * solo.std
                 and abs(solo.tail_sum - shard.tail_sum) <= 1e-12 * abs(solo.tail_sum))
        report[name] = dict(world=world, identical_on_every_rank=same_on_all_ranks, order_statistics_and_counts_equal_one_gpu=bool(exact),
                            moments_equal_to_rounding=bool(close), p50=shard.pct.get("p50"))
    dist.barrier()
if rank == 0:
    report["_exchange"] = dict(peer_memory_attached=bool(getattr(eng, "_peer_state", (0, False))[1]),
                               fused_runs=eng.debug_counter(0), fused_runs_redone_by_staged_selection=eng.debug_counter(1))
    print(json.dumps(report, indent=1))
    ok = all(v["identical_on_every_rank"] and v["order_statistics_and_counts_equal_one_gpu"] and v["moments_equal_to_rounding"]
             for k, v in report.items() if not k.startswith("_"))
    print("NCCL PARITY", "OK" if ok else "FAILED")
dist.destroy_process_group()
