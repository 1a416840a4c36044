# Round-end evidence in one call: GPU tests, smoke, default bench + reference arm (also for c1), launch lists and
# ncu --set full captures of the dominant kernels.  Outputs land in gpurun_out/ and are copied to profiles/r02/ by hand.
set -x
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 1500 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo default rc=$?
timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo reference rc=$?
timeout 900 python bench.py --impl reference --workload c1 > gpurun_out/bench_reference_c1.json 2> gpurun_out/bench_reference_c1.err; echo reference c1 rc=$?
for wl in c2 ref c5 c3 ref64; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --workload $wl > gpurun_out/plain_$wl.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$wl.csv \
      python bench.py --steps 2 --warmup 3 --no-extras --workload $wl > gpurun_out/ncu_$wl.log 2>&1
  echo ncu $wl rc=$?
done
python tools/timeline.py c2 2>&1 | grep -v -i warn > gpurun_out/timeline_c2.txt
python tools/timeline.py c5 2>&1 | grep -v -i warn > gpurun_out/timeline_c5.txt
python tools/timeline.py ref 2>&1 | grep -v -i warn > gpurun_out/timeline_ref.txt
python tools/c1_latency.py > gpurun_out/c1_latency.txt 2>&1
bash tools/gpu_ncu_full.sh "c2 sim32_kernel simc2 3" "ref sim32_kernel simref 3" "c2 fused_closing k4 3"
