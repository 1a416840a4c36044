set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -5 gpurun_out/pytest.log
for wl in c2 c5; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1
  echo bench $wl rc=$?; tail -1 gpurun_out/plain_$wl.log | cut -c1-330
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/ncu_$wl.log 2>&1
done
