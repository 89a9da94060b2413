timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bash tools/gpu_ab.sh cur 2>&1 | grep "^AB"
BENCH_ARGS="--variant B" bash tools/gpu_ab.sh cur 2>&1 | grep "^AB" | sed 's/^AB/AB-varB/'
BENCH_ARGS="--workload rollout --variant B" bash tools/gpu_ab.sh cur 2>&1 | grep "^AB" | sed 's/^AB/AB-cfg3/'
BENCH_ARGS="--workload cfg5" STEPS=3 bash tools/gpu_ab.sh cur 2>&1 | grep "^AB" | sed 's/^AB/AB-cfg5/'
timeout 100 python tools/phase_cycles.py > gpurun_out/phase_dev.json 2>&1; python -c "
import json; d=json.load(open('gpurun_out/phase_dev.json')); print(d['ms_profiled_launch'], d['phases_pct_of_warp_time'], d['epochs'], d['replays'])"
