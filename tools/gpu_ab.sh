# same-box A/B of renderer builds / options:  tools/gpu_ab.sh "<tag>[:opt=val,...]" ...   (tag "" = the in-tree library; others from tools/ab_build.sh)
set -x
B="--steps ${STEPS:-12} --warmup 3 --no-cpu-baseline --no-ref-gpu ${BENCH_ARGS:-}"
for spec in "$@"; do
  tag=${spec%%:*}; opts=; [ "$spec" != "$tag" ] && opts=$(echo "${spec#*:}" | tr ',' ' ' | sed 's/[^ ]* */--opt &/g')
  [ "$tag" = "cur" ] && tag=
  lib=; [ -n "$tag" ] && lib=$PWD/nerfsafetyvalidation_b200/_lib/ab/libngp_$tag.so
  out=gpurun_out/ab_$(echo "${spec:-cur}" | tr ':=,' '___')
  NGP_B200_LIB=$lib timeout 200 python bench.py $B $opts > $out.json 2> $out.err
  python -c "
import json,sys; d=json.load(open('$out.json')); print('AB ${spec:-cur}', round(d['ms_per_step'],3), 'ms; epochs', d['config'].get('epochs_per_step'), 'replays', d['config'].get('replayed_epochs'))" || tail -3 $out.err
done
