timeout 900 python -m pytest tests/test_gpu_api.py -q -x -k "regeneration or fan_chart or without_paths or shard" 2>&1 | tail -3
timeout 300 python bench.py --steps 3 --warmup 2 --no-extras --workload c4fan 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('c4fan 1 gpu ms', round(d['ms_per_step'],3))"
