set -x
timeout 600 python -m pytest tests/test_gpu_fused.py tests/test_zz_no_alloc.py -q -x 2>&1 | tail -3
python tools/timeline.py c2 2>&1 | grep -v -i warn | tail -6
