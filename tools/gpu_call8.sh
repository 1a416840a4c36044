set -x
S=$(date +%s); timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo default rc=$? wall=$(( $(date +%s) - S ))s
tail -1 gpurun_out/bench_default.json | cut -c1-1800
S=$(date +%s); timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo reference rc=$? wall=$(( $(date +%s) - S ))s
tail -1 gpurun_out/bench_reference.json | cut -c1-800
