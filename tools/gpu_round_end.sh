# What the driver runs at round end, in one gpurun call: GPU tests, smoke, default bench + reference arm,
# and the ncu launch lists of the bench workloads (plain run first, each only after exit 0).
set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo default rc=$?
timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo reference rc=$?
for wl in "$@"; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/ncu_$wl.log 2>&1
  echo ncu $wl rc=$?
done
