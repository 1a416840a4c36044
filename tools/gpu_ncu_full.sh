set -x
for spec in "c2 sim32_kernel simc2" "c5 sim32_kernel simc5" "c5 select_pass1_f32_kernel<5> pass1c5" "c5 select_pass2_f32_kernel pass2c5" "ref sim32_kernel simref" "c3 portfolio_kernel simc3"; do
  set -- $spec
  timeout 300 python bench.py --steps 1 --warmup 3 --no-extras --workload $1 > gpurun_out/plain_full_$3.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$2" -s 4 -c 1 -f -o gpurun_out/full_$3 \
      python bench.py --steps 1 --warmup 3 --no-extras --workload $1 > gpurun_out/ncu_full_$3.log 2>&1
  echo full $3 rc=$?
done
ls -la gpurun_out/*.ncu-rep
