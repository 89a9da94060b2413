set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gpu_tests_k.log
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err
timeout 300 python tools/phase_cycles.py > gpurun_out/phase_k.json 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_k.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_k1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k k_render -c 1 -s 2 -o gpurun_out/prof_r01k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_k2.log 2>&1
cat gpurun_out/gpu_tests_k.log; cat gpurun_out/bench_k.json
