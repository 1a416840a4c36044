set -x
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -6 gpurun_out/pytest.log
python tools/timeline.py c5 2>&1 | grep -v Warn | tail -40 | cut -c1-130
for wl in c5 c4fan; do
timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>gpurun_out/bench_$wl.err | tail -1 > gpurun_out/bench_$wl.json; python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl', 'ms', round(d['ms_per_step'],4), 'sim', round(d['sim_kernel_ms'],4), 'e2e ms', round(d['e2e']['ms_per_step'],4), 'frac', d['roofline']['frac'])"
done
