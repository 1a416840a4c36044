set -x
timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload c2 > gpurun_out/plain_c2.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file gpurun_out/launches_c2_warm.csv \
    python bench.py --steps 2 --warmup 3 --no-extras --workload c2 > gpurun_out/ncu_c2_warm.log 2>&1
echo rc=$?
