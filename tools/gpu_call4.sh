set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -30 gpurun_out/pytest.log
for wl in c2 c5; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1
  echo bench $wl rc=$?; tail -3 gpurun_out/plain_$wl.log
done
