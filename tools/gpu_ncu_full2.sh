set -x
for spec in "c5 select_pass1_f32_kernel pass1c5 5" "c5 select_pass2_f32_kernel pass2c5 5"; do
  set -- $spec
  timeout 300 python bench.py --steps 1 --warmup 3 --no-extras --workload $1 > gpurun_out/plain_full_$3.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$2" -s $4 -c 1 -f -o gpurun_out/full_$3 \
      python bench.py --steps 1 --warmup 3 --no-extras --workload $1 > gpurun_out/ncu_full_$3.log 2>&1
  echo full $3 rc=$?
done
