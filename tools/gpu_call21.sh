timeout 900 python -m pytest tests/test_gpu_api.py tests/test_gpu_rng_select.py tests/test_gpu_select_stress.py -q -x 2>&1 | tail -1
python tools/timeline.py c5 2>&1 | grep -v -i warn | grep "pass2\|span" | cut -c1-120
for wl in c5 c4fan; do timeout 300 python bench.py --steps 8 --warmup 3 --no-extras --workload $wl 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$wl ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"; done
