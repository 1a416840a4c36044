set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -5          # alone, fresh process
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo full rc=$?; tail -6 gpurun_out/pytest.log
