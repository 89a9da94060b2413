# ablation builds: bench only (their results are wrong by construction)
for tag in $TAGS; do
  export NGP_B200_LIB=$PWD/nerfsafetyvalidation_b200/_lib/ab/libngp_$tag.so
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_$tag.json')); print('MS_PER_STEP $tag', d['ms_per_step'], 'samples', d['config']['samples_per_step'], 'iters', d['config']['iterations_per_step'])"
done
