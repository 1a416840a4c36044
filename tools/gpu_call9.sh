set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fused.py -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -4 gpurun_out/pytest.log
for wl in ref c1; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --workload $wl 2>gpurun_out/bench_$wl.err | tail -1 > gpurun_out/bench_$wl.json; python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl', 'ms', round(d['ms_per_step'],4), 'sim', round(d['sim_kernel_ms'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4), 'frac', d['roofline']['frac'], d['roofline']['survey_8d']['frac'])"
done
