# usage: bash tools/gpu_ncu_full.sh "<workload> <kernel regex> <tag> <skip>" ...   (<= 3 captures per call: the reports are ~13 MB each
# and gpurun brings back at most 64 MiB)
set -x
for spec in "$@"; do
  set -- $spec
  timeout 300 python bench.py --steps 1 --warmup 3 --no-extras --workload $1 > gpurun_out/plain_full_$3.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$2" -s $4 -c 1 -f -o gpurun_out/full_$3 \
      python bench.py --steps 1 --warmup 3 --no-extras --workload $1 > gpurun_out/ncu_full_$3.log 2>&1
  echo full $3 rc=$?
done
ls -la gpurun_out/*.ncu-rep
