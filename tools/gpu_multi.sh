# multi-GPU check: bench.py under torchrun at N = $NGPU, both arms, plus the rollout / CEM stress tool
N=${NGPU:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo rc=$?
tail -3 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/stress_test.py --mode cem --iters 3 --rollouts 256 > gpurun_out/stress_n$N.log 2>&1; echo rc=$?
tail -12 gpurun_out/stress_n$N.log
