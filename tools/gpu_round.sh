# the round's evidence set: GPU tests, bench lines, launch list, --set full captures of k_render per workload, phase cycles, ablation
set -x
TAG=${1:-r02h}
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -5 > gpurun_out/${TAG}_gpu_tests.log; cat gpurun_out/${TAG}_gpu_tests.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench_cfg2.json 2> gpurun_out/bench.err || tail -5 gpurun_out/bench.err
timeout 300 python bench.py --variant B --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg2_variantB.json 2>> gpurun_out/bench.err
timeout 300 python bench.py --workload rollout --variant B --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_rollout_cfg3batch_B.json 2>> gpurun_out/bench.err
timeout 400 python bench.py --workload cfg5 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg5_1gpu.json 2>> gpurun_out/bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_l.log 2>&1
for wl in cfg2 rollout cfg5; do
  extra=; [ $wl = rollout ] && extra="--variant B"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -s 2 -o gpurun_out/prof_${TAG}_$wl -f python bench.py --workload $wl $extra --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_$wl.log 2>&1; tail -1 gpurun_out/ncu_$wl.log
done
timeout 100 python tools/phase_cycles.py > gpurun_out/${TAG}_phase_cycles.json 2>&1
for tag in nogather nomlp noboth; do
  NGP_B200_LIB=$PWD/nerfsafetyvalidation_b200/_lib/ab/libngp_$tag.so timeout 200 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/abl_$tag.json 2> gpurun_out/abl_$tag.err
done
python - <<PY
import json
out={"what":"cfg2 frame, same box, same run: full kernel vs builds without the hash-grid gather (operand = zeros), without the MLPs (no MMAs / epilogues), without both (tools/ab_build.sh -DNGP_ABLATE_*)"}
out["full_ms"]=json.load(open("gpurun_out/${TAG}_bench_cfg2.json"))["ms_per_step"]
for t in ("nogather","nomlp","noboth"):
    try: out[t+"_ms"]=json.load(open(f"gpurun_out/abl_{t}.json"))["ms_per_step"]
    except Exception as e: out[t+"_ms"]=None
json.dump(out,open("gpurun_out/${TAG}_ablation_cfg2.json","w"),indent=1); print(out)
PY
