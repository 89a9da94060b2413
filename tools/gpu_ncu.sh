# one --set full capture of k_render on the bench frame (+ the launch list), after the plain run has succeeded
set -x
TAG=${1:-r02}
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/bench_pre_ncu.json 2> gpurun_out/bench_pre_ncu.err || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -s 2 -o gpurun_out/prof_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_full.log 2>&1; tail -3 gpurun_out/ncu_full.log
