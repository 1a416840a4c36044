timeout 900 python -m pytest tests/test_gpu_rng_select.py tests/test_gpu_select_stress.py tests/test_gpu_api.py -q -x 2>&1 | tail -2
python tools/timeline.py c5 2>&1 | grep -v -i warn | grep "bracket\|span"
for wl in c5; do timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$wl ms', round(d['ms_per_step'],3))"; done
