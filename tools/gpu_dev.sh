set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
bash tools/gpu_scale_stress.sh 1 2>&1 | grep -v "^+"
