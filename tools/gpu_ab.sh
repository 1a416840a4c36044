# GPU tests + c5 / c4fan bench lines + launch list of c5 + full capture of one kernel
# usage: bash tools/gpu_ab.sh [kernel-regex] [ENV=1 ...]   (extra env settings get their own c5 line)
set -x
timeout 700 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest.log
for wl in c5 c4fan; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$wl', round(d['ms_per_step'],3), d['result'] if 'result' in d else '')"
done
K=$1; shift
for e in "$@"; do
  env $e timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload c5 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('c5 $e', round(d['ms_per_step'],3))"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c5.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload c5 > gpurun_out/ncu_c5.log 2>&1
echo ncu rc=$?
if [ -n "$K" ]; then bash tools/gpu_ncu_full.sh "c5 $K k 2"; fi
