# multi-GPU check after a kernel change: NCCL parity (global job on W GPUs = the one-GPU numbers) + bench lines
# usage (inside gpurun --gpus W): bash tools/gpu_multi.sh W
W=$1
set -x
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 tools/nccl_parity.py > gpurun_out/nccl_parity_${W}gpu.txt 2> gpurun_out/nccl_parity_${W}gpu.err; echo parity rc=$?
for wl in ${WLS:-c2 c5 c4fan}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $W --steps 5 --warmup 3 --no-extras --workload $wl 2> gpurun_out/bench_${W}gpu_$wl.err | tail -1 >> gpurun_out/bench_${W}gpu.txt; echo $wl rc=$?
done
