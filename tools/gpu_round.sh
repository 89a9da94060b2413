set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gpu_tests_n.log
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n.json 2> gpurun_out/bench_n.err
timeout 300 python tools/phase_cycles.py > gpurun_out/phase_n.json 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_n.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_n1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k k_render -c 1 -s 2 -o gpurun_out/prof_r01n -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/ncu_n2.log 2>&1
cat gpurun_out/gpu_tests_n.log; cat gpurun_out/bench_n.json
