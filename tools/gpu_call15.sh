timeout 600 python -m pytest tests/test_gpu_fused.py -q -x 2>&1 | tail -2
python tools/timeline.py c2 2>&1 | grep -v -i warn | tail -5
run() { MC_BENCH_SINK=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-extras --workload $1 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$1 sink=$2', 'ms', round(d['ms_per_step'],4), 'sim', round(d['sim_kernel_ms'],4), 'p50' , d['result'].get('var_95'))"; }
run c2 0,0,3; run c2 1,0,3; run c2 2,0,3
run c4 0,0,3; run c4 1,0,3
run ref 0,0,3; run ref 0,1,2; run ref 0,2,2; run ref 0,1,1; run ref 0,0,2
