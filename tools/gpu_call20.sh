timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo full rc=$?; tail -4 gpurun_out/pytest.log
python tools/timeline.py c5 2>&1 | grep -v -i warn | grep "bracket\|pass2\|pass1\|span"| cut -c1-110
for wl in c5 c5 c4fan; do timeout 300 python bench.py --steps 8 --warmup 3 --no-extras --workload $wl 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$wl ms', round(d['ms_per_step'],3))"; done
