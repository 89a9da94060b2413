# development iteration on the GPU box: tcgen05 network kernel vs mma.sync, the GPU test suite, a short bench, the phase profile
set -x
timeout 200 python tools/tc_check.py > gpurun_out/tc_check.json 2> gpurun_out/tc_check.err; python -c "
import json; d=json.load(open('gpurun_out/tc_check.json')); print('TC_CHECK max sigma_rel', max(v['sigma_rel'] for v in d.values()), 'max rgb_abs', max(v['rgb_abs'] for v in d.values()), 'ms 1M', d['A/1048576']['ms_mma_sync'], d['A/1048576']['ms_tcgen05'])"; tail -3 gpurun_out/tc_check.err
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/gpu_tests_dev.log; cat gpurun_out/gpu_tests_dev.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_dev.json 2> gpurun_out/bench_dev.err; tail -c 600 gpurun_out/bench_dev.err
python -c "
import json; d=json.load(open('gpurun_out/bench_dev.json')); print('MS_PER_STEP', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], d['config'].get('epochs_per_step'), d['config'].get('replayed_epochs'))"
timeout 100 python tools/phase_cycles.py > gpurun_out/phase_dev.json 2>&1; python -c "
import json; d=json.load(open('gpurun_out/phase_dev.json')); print(d['ms_profiled_launch'], d['phases_pct_of_warp_time'], d['epochs'], d['replays'])"
