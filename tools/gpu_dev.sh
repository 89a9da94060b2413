set -x
timeout 900 python -m pytest tests/test_gpu_occupancy.py tests/test_gpu_uq.py -x -q 2>&1 | tail -30
