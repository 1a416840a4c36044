set -x
timeout 600 python -m pytest tests/test_gpu_fused.py -q -x > gpurun_out/pytest_fused.log 2>&1; echo fused rc=$?; tail -25 gpurun_out/pytest_fused.log
timeout 900 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_fused.py > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -15 gpurun_out/pytest.log
for wl in c2 c4 ref c3; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --workload $wl 2>gpurun_out/bench_$wl.err | tail -1 > gpurun_out/bench_$wl.json; python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl', 'ms', round(d['ms_per_step'],4), 'sim', round(d['sim_kernel_ms'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4), 'launches', d['gpu_launches'], d['result'])"
done
