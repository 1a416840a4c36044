set -x
timeout 900 python -m pytest tests/test_gpu_normals.py tests/test_gpu_api.py::test_statistical_parity_with_reference_c1 -q -x > gpurun_out/pytest_normals.log 2>&1; echo normals rc=$?; tail -15 gpurun_out/pytest_normals.log
for wl in c2full c2; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --workload $wl 2>gpurun_out/bench_$wl.err | tail -1 > gpurun_out/bench_$wl.json; python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl', 'ms', round(d['ms_per_step'],4), 'sim', round(d['sim_kernel_ms'],4), 'frac', d['roofline']['frac'], d['roofline']['survey_8d']['frac'])"
done
