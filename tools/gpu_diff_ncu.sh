# evidence of the differentiable small-batch path: its tests' printed gradient errors, the latency line, the ncu launch
# list of one forward + backward, and one --set full capture of each of its kernels -- each ncu pass only after the plain run succeeded
set -x
TAG=${1:-r02i}
timeout 500 python -m pytest tests/test_gpu_diff.py -q -s > gpurun_out/${TAG}_diff_tests.log 2>&1 || exit 1
timeout 300 python tools/diff_bench.py > gpurun_out/${TAG}_diff_bench.json 2> gpurun_out/diff_bench.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_diff_launches.csv \
    python tools/diff_bench.py --iters 3 --warmup 1 > gpurun_out/ncu_diff_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_diff_render|k_density_backward' -c 3 -s 6 -o gpurun_out/prof_${TAG}_diff -f \
    python tools/diff_bench.py --iters 3 --warmup 1 > gpurun_out/ncu_diff_full.log 2>&1; tail -3 gpurun_out/ncu_diff_full.log
timeout 300 python tools/small_batch_probe.py > gpurun_out/${TAG}_small_batch.json 2> gpurun_out/small_batch.err
