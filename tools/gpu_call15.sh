for wl in c2 c4 c4fan; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 5 --warmup 3 --workload $wl --no-extras 2> gpurun_out/bench8_$wl.err | tail -1 > gpurun_out/bench8_$wl.json
  echo "$wl rc=$?"; python - <<PY
import json
d = json.load(open("gpurun_out/bench8_$wl.json"))
print("$wl", {k: d[k] for k in ("value", "ms_per_step", "n_gpus", "sim_kernel_ms", "gpu_launches")}, d["e2e"]["value"], d["clocks"])
PY
done
