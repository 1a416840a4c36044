# compute-sanitizer over the reduction / selection / histogram kernels: memcheck, then racecheck on a
# smaller subset (racecheck instruments every shared-memory access).  Summaries land in gpurun_out/.
set -x
SEL="tests/test_gpu_rng_select.py tests/test_gpu_parity.py::test_device_bootstrap_run_matches_port tests/test_gpu_parity.py::test_gbm_injected_and_device_normals tests/test_gpu_parity.py::test_portfolio_matches_port tests/test_gpu_api.py::test_shard_invariance_emulated_ranks"
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20 --log-file gpurun_out/sanitize_memcheck.log \
    python -m pytest $SEL tests/test_gpu_select_stress.py -q -x -k "not 400" > gpurun_out/sanitize_memcheck_pytest.log 2>&1
echo memcheck rc=$?; tail -2 gpurun_out/sanitize_memcheck_pytest.log; tail -3 gpurun_out/sanitize_memcheck.log
timeout 1500 compute-sanitizer --tool racecheck --racecheck-report analysis --error-exitcode 9 --print-limit 20 --log-file gpurun_out/sanitize_racecheck.log \
    python -m pytest tests/test_gpu_rng_select.py "tests/test_gpu_parity.py::test_device_bootstrap_run_matches_port" -q -x > gpurun_out/sanitize_racecheck_pytest.log 2>&1
echo racecheck rc=$?; tail -2 gpurun_out/sanitize_racecheck_pytest.log; tail -3 gpurun_out/sanitize_racecheck.log
