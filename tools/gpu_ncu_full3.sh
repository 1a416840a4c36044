set -x
timeout 300 python bench.py --steps 1 --warmup 3 --no-extras --workload c2 > gpurun_out/plain_full_c2close.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on \
    -k "regex:select_pass1_f32_kernel|select_pass2_f32_kernel|tail_hist_kernel|select_bracket_kernel|reduce_partials_kernel" -s 15 -c 5 -f -o gpurun_out/full_c2close \
    python bench.py --steps 1 --warmup 3 --no-extras --workload c2 > gpurun_out/ncu_full_c2close.log 2>&1
echo rc=$?
