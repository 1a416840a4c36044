for nb in 256 512; do for w in 2 4 8 12 24; do
  for wl in c2 c4; do
  MC_BENCH_NBLOCK=$nb MC_BENCH_WAVES=$w timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$wl nb=$nb waves=$w', round(d['ms_per_step'],4), round(d['sim_kernel_ms'],4))"
  done
done; done
