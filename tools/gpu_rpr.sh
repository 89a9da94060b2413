for r in 32 64; do
  for args in "--workload cfg2" "--workload cfg2 --variant B" "--workload rollout --variant B" "--workload cfg5"; do
    NGP_ROWS_PER_RAY=$r timeout 300 python bench.py $args --steps 8 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/r0.json 2> gpurun_out/r0.err
    python -c "
import json; d=json.load(open('gpurun_out/r0.json')); c=d['config']; print('RPR $r $args | ms', round(d['ms_per_step'],3), 'iters', c['iterations_per_step'], 'epochs', c['epochs_per_step'], 'replays', c['replayed_epochs'])"
  done
done
