# the multi-GPU lines of the final build in one pass, cheapest first:  gpurun --gpus N -- 'bash tools/gpu_scale_quick.sh N TAG'
N=${1:-8}; TAG=${2:-r02j}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 200 $RUN bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/${TAG}_scale_cfg2_n$N.json 2> gpurun_out/scale_cfg2_n$N.err
timeout 200 $RUN bench.py --gpus $N --workload cfg5 --split-rows --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/${TAG}_scale_cfg5_rows_n$N.json 2> gpurun_out/scale_cfg5_n$N.err
timeout 200 $RUN tools/stress_test.py --mode mc --rollouts 4096 --steps 20 --res 200 > gpurun_out/${TAG}_scale_cfg3_mc_n$N.json 2> gpurun_out/scale_mc_n$N.err
timeout 300 $RUN tools/stress_test.py --mode cem --iters 8 --rollouts 2048 --steps 20 --res 200 > gpurun_out/${TAG}_scale_cfg4_cem_n$N.json 2> gpurun_out/scale_cem_n$N.err
for f in cfg2 cfg5_rows cfg3_mc cfg4_cem; do tail -c 400 gpurun_out/${TAG}_scale_${f}_n$N.json; echo; done
