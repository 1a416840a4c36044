MC_BENCH_DEBUG=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload c5 2>&1 >/dev/null | grep per-step
MC_BENCH_DEBUG=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload c2 2>&1 >/dev/null | grep per-step
MC_BENCH_DEBUG=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-extras --workload c4fan 2>&1 >/dev/null | grep per-step
