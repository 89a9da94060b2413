# quick GPU iteration: parity tests of the renderer, then a short bench and the phase profile
set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err
python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print('MS_PER_STEP', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'])"
timeout 300 python tools/phase_cycles.py > gpurun_out/phase_q.json 2>&1; python -c "
import json; d=json.load(open('gpurun_out/phase_q.json')); print(d['phases_pct_of_warp_time'], d['epochs'], d['replays'])"
