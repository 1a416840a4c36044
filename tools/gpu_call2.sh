set -x
./tools/microbench > gpurun_out/microbench.txt 2>&1; echo microbench rc=$?
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -15 gpurun_out/pytest.log
for wl in c2 ref c5; do
  python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/ncu_$wl.log 2>&1
  echo ncu-launches $wl rc=$?
done
python bench.py --steps 2 --warmup 3 --no-extras --workload c2 > gpurun_out/plain_c2b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sim_kernel -s 4 -c 1 -f -o gpurun_out/prof_c2 \
    python bench.py --steps 2 --warmup 3 --no-extras --workload c2 > gpurun_out/ncu_full_c2.log 2>&1; echo ncu-full c2 rc=$?
python bench.py --steps 2 --warmup 3 --no-extras --workload ref > gpurun_out/plain_refb.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sim_kernel -s 4 -c 1 -f -o gpurun_out/prof_ref \
    python bench.py --steps 2 --warmup 3 --no-extras --workload ref > gpurun_out/ncu_full_ref.log 2>&1; echo ncu-full ref rc=$?
ls -la gpurun_out
