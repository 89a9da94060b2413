set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_dev.json 2> gpurun_out/bench_dev.err; tail -c 600 gpurun_out/bench_dev.err
python -c "
import json; d=json.load(open('gpurun_out/bench_dev.json')); print('MS_PER_STEP', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'pose', d['e2e_from_pose']['value'], d['e2e_from_pose']['ms_per_step'])"
timeout 300 python bench.py --workload rollout --variant B --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_dev3.json 2> gpurun_out/bench_dev3.err
python -c "
import json; d=json.load(open('gpurun_out/bench_dev3.json')); print('CFG3 MS_PER_STEP', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'pose', d['e2e_from_pose']['value'], d['e2e_from_pose']['ms_per_step'])"
