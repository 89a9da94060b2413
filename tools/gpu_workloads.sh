# the other BASELINE.json configs as bench workloads (1 GPU): cfg5-shaped frame and the cfg3 rollout batch, variant B too
for args in "--workload cfg5" "--workload rollout --variant B" "--workload cfg2 --variant B"; do
  tag=$(echo $args | tr -d ' -')
  timeout 600 python bench.py $args --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "rc=$? $args"
  tail -2 gpurun_out/bench_$tag.err; python -c "
import json; d=json.load(open('gpurun_out/bench_$tag.json')); print(d['config']['workload'], '| ms', d['ms_per_step'], 'Mrays/s', d['value'], 'e2e', d['e2e']['value'], 'samples/ray', d['config']['samples_per_ray'], 'iters', d['config']['iterations_per_step'], 'ref_gpu', d.get('reference_gpu'))"
done
