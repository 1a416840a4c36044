set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo pytest rc=$?; tail -4 gpurun_out/pytest.log
python tools/host_overhead.py c2 2>&1 | head -40
timeout 1200 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo default rc=$?; tail -3 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo reference rc=$?
