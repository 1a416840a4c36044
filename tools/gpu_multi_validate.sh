# N-GPU validation (inside gpurun --gpus W): NCCL parity of the global jobs, default bench with the multi-rank extras
# (c4, c4fan), device timeline of one sharded c2 job.   usage: bash tools/gpu_multi_validate.sh W
set -x
W=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 tools/nccl_parity.py > gpurun_out/nccl_parity_${W}gpu.txt 2> gpurun_out/nccl_parity_${W}gpu.err; echo parity rc=$?; tail -9 gpurun_out/nccl_parity_${W}gpu.txt; grep -c true gpurun_out/nccl_parity_${W}gpu.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $W --steps 10 --warmup 3 2> gpurun_out/bench_${W}gpu.err | tail -1 > gpurun_out/bench_${W}gpu.json; tail -2 gpurun_out/bench_${W}gpu.err
python -c "
import json
d=json.load(open('gpurun_out/bench_${W}gpu.json'))
print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['ms_per_step'])
for k,v in d['extras'].items(): print(k, v.get('ms_per_step'), v.get('value'), v.get('error'))"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29513 tools/timeline_sharded.py c2 2>&1 | grep -v Warn | tail -6 | tee gpurun_out/timeline_sharded_${W}gpu_c2.txt
