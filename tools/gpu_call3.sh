set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -8 gpurun_out/pytest.log
timeout 600 python bench.py > gpurun_out/bench_default.log 2>&1; echo bench rc=$?; tail -2 gpurun_out/bench_default.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo benchref rc=$?; tail -1 gpurun_out/bench_ref.log
for wl in c2 ref c5 c3; do
  timeout 300 python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/ncu_$wl.log 2>&1
  echo ncu-launches $wl rc=$?; tail -1 gpurun_out/plain_$wl.log
done
ls -la gpurun_out
