set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -5 gpurun_out/pytest.log
for wl in c5 c2 c4fan; do
  MC_BENCH_DEBUG=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>&1 >/dev/null | grep per-step
  MC_NO_SIDE_LISTS=1 MC_BENCH_DEBUG=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>&1 >/dev/null | grep per-step
done
