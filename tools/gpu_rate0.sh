for args in "--workload cfg2" "--workload cfg2 --variant B" "--workload rollout --variant B" "--workload cfg5"; do
    timeout 300 python bench.py $args --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/r0.json 2> gpurun_out/r0.err
    python -c "
import json; d=json.load(open('gpurun_out/r0.json')); c=d['config']; print('OPAQUE', '$args', '| ms', round(d['ms_per_step'],3), 'iters', c['iterations_per_step'], 'epochs', c['epochs_per_step'], 'replays', c['replayed_epochs'])"
done
timeout 600 python -m pytest tests -m gpu -x -q -k "render or cfg" 2>&1 | tail -3
