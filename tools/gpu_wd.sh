set -x
B="--steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu"
for tag in "" noregs; do
  lib=; [ -n "$tag" ] && lib=$PWD/nerfsafetyvalidation_b200/_lib/ab/libngp_$tag.so
  NGP_B200_LIB=$lib timeout 200 python bench.py $B > gpurun_out/abl_${tag:-full}.json 2> gpurun_out/abl_${tag:-full}.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/abl_${tag:-full}.json')); print('ABLATE ${tag:-full}', d['ms_per_step'])"
  NGP_B200_LIB=$lib timeout 60 python -X faulthandler tools/phase_cycles.py > gpurun_out/phase_${tag:-full}.json 2> gpurun_out/phase_${tag:-full}.err; echo "phase ${tag:-full} rc=$?"; head -c 2500 gpurun_out/phase_${tag:-full}.json; tail -3 gpurun_out/phase_${tag:-full}.err
done
