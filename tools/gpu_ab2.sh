bash tools/gpu_ab.sh base cur 2>&1 | grep "^AB"
BENCH_ARGS="--variant B" bash tools/gpu_ab.sh cur 2>&1 | grep "^AB" | sed 's/^AB/AB-varB/'
BENCH_ARGS="--workload rollout --variant B" bash tools/gpu_ab.sh cur 2>&1 | grep "^AB" | sed 's/^AB/AB-cfg3/'
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python tools/parity_margin.py 2>&1 | tail -20
