timeout 900 python -m pytest tests/test_gpu_api.py tests/test_caller_shim.py tests/test_inputs.py tests/test_gpu_parity.py -q -x 2>&1 | tail -2
python tools/small_call_breakdown.py 2>&1 | grep " us "
python tools/c1_latency.py 2>&1 | head -3
