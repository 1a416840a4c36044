set -x
timeout 600 python -m pytest tests/test_gpu_fused.py -q -x > gpurun_out/pytest_fused.log 2>&1; echo fused rc=$?; tail -3 gpurun_out/pytest_fused.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/timeline_sharded.py c2 2>&1 | grep -v Warn | tail -6
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras 2> gpurun_out/bench_2gpu.err | tail -1 > gpurun_out/bench_2gpu.json
python -c "
import json
d=json.load(open('gpurun_out/bench_2gpu.json'))
print(d['n_gpus'], d['ms_per_step'], d['value'], d['e2e']['ms_per_step'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/nccl_parity.py > gpurun_out/nccl_parity_2gpu.txt 2> gpurun_out/nccl_parity_2gpu.err; tail -8 gpurun_out/nccl_parity_2gpu.txt
