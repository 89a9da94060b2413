# what does the frame cost without one of the two big phases?  (libraries from tools/ab_build.sh <tag> -DNGP_ABLATE_...)
set -x
B="--steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu"
for tag in "" nogather nomlp noboth; do
  lib=; [ -n "$tag" ] && lib=$PWD/nerfsafetyvalidation_b200/_lib/ab/libngp_$tag.so
  NGP_B200_LIB=$lib timeout 200 python bench.py $B > gpurun_out/abl_${tag:-full}.json 2> gpurun_out/abl_${tag:-full}.err
  python -c "
import json,sys; d=json.load(open('gpurun_out/abl_${tag:-full}.json')); print('ABLATE ${tag:-full}', d['ms_per_step'])"
done
timeout 120 python -X faulthandler tools/phase_cycles.py > gpurun_out/phase_dev.json 2> gpurun_out/phase_dev.err; echo "phase rc=$?"; tail -5 gpurun_out/phase_dev.err; cat gpurun_out/phase_dev.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -s 2 -o gpurun_out/prof_r02b -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_full.log 2>&1; tail -3 gpurun_out/ncu_full.log
