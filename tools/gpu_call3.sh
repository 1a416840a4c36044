set -x
timeout 600 python -m pytest tests/test_gpu_fused.py -q -x > gpurun_out/pytest_fused.log 2>&1; echo fused rc=$?; tail -5 gpurun_out/pytest_fused.log
python tools/timeline.py c2 2>&1 | tail -7
python tools/timeline.py ref 2>&1 | tail -7
python tools/timeline.py c4 2>&1 | tail -7
for wl in c2 c4 ref; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --workload $wl 2>gpurun_out/bench_$wl.err | tail -1 > gpurun_out/bench_$wl.json; python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl', 'ms', round(d['ms_per_step'],4), 'sim', round(d['sim_kernel_ms'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4), 'launches', d['gpu_launches'])"
done
