timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo default rc=$?
timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo reference rc=$?
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_default.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "sim_kernel_ms", "gpu_launches")}, d["e2e"], d["clocks"], d["roofline"]["frac"], d["cpu_baseline"])
for k, v in d["extras"].items():
    print(k, {a: v.get(a) for a in ("value", "ms_per_step", "sim_kernel_ms", "error")} if k != "callers" else v)
PY
