for args in "--workload rollout --variant B" "--workload cfg2"; do
  timeout 300 python bench.py $args --steps 8 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/r0.json 2> gpurun_out/r0.err; tail -2 gpurun_out/r0.err
  python -c "
import json; d=json.load(open('gpurun_out/r0.json')); print('$args | dev', round(d['value'],2), 'e2e rays', round(d['e2e']['value'],2), d['e2e']['h2d_bytes_per_step'], 'e2e pose', round(d['e2e_from_pose']['value'],2), d['e2e_from_pose']['h2d_bytes_per_step'])"
done
