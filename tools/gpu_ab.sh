# A/B of renderer builds: for each tag in $TAGS run the renderer parity tests and a short bench
for tag in $TAGS; do
  export NGP_B200_LIB=$PWD/nerfsafetyvalidation_b200/_lib/ab/libngp_$tag.so
  [ "$tag" = base ] && unset NGP_B200_LIB
  echo "=== $tag"
  timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render" 2>&1 | tail -3
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_$tag.json')); print('MS_PER_STEP $tag', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'])"
done
