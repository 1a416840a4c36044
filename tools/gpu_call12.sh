set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -5 gpurun_out/pytest.log
for wl in c4fan c5; do
  MC_BENCH_DEBUG=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>&1 >/dev/null | grep per-step
done
