set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -5 gpurun_out/pytest.log
for wl in c2 ref c4 c3 c5; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1
  echo bench $wl rc=$?; tail -1 gpurun_out/plain_$wl.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: round(d[k],4) if isinstance(d[k], float) else d[k] for k in ('value','ms_per_step','sim_kernel_ms','gpu_launches')}, round(d['e2e']['value']/1e9,1), round(d['roofline']['frac'],4))"
done
