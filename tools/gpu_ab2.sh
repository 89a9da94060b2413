bash tools/gpu_ab.sh cur cur:rows_per_ray=32 2>&1 | grep "^AB"
BENCH_ARGS="--variant B" bash tools/gpu_ab.sh cur cur:rows_per_ray=32  2>&1 | grep "^AB" | sed 's/^AB/AB-varB/'
BENCH_ARGS="--workload rollout --variant B" bash tools/gpu_ab.sh cur cur:rows_per_ray=32 2>&1 | grep "^AB" | sed 's/^AB/AB-cfg3/'
BENCH_ARGS="--workload cfg5" STEPS=3 bash tools/gpu_ab.sh cur cur:rows_per_ray=32 2>&1 | grep "^AB" | sed 's/^AB/AB-cfg5/'
