set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -30 gpurun_out/pytest.log
for wl in c2 c5 ref c3 c4; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1
  echo bench $wl rc=$?; tail -1 gpurun_out/plain_$wl.log | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d[k] for k in ('value','ms_per_step','sim_kernel_ms','gpu_launches')}, d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'])"
done
