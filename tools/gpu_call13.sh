set -x
for wl in c5; do
  timeout 300 python bench.py --steps 3 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/ncu_$wl.log 2>&1
  echo ncu $wl rc=$?
done
