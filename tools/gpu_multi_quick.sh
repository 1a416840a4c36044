# quick N-GPU look at the sharded fused path: device timeline of a c2 job, c2 and c4 bench lines without extras
W=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29513 tools/timeline_sharded.py c2 2>&1 | grep -v Warn | tail -6
for wl in c2 c4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $W --steps 10 --warmup 3 --no-extras --workload $wl 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$wl', d['n_gpus'], 'ms', round(d['ms_per_step'],4), 'value %.4e' % d['value'], 'e2e ms', round(d['e2e']['ms_per_step'],4))"
done
