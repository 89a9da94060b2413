# scaling record: bench.py at N = 1, 2, 4, 8 on one box (cfg2), then the cfg5 frame and the CEM stress shape on all 8 GPUs
for N in 1 2 4 8; do
  if [ $N = 1 ]; then timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; fi
  echo "N=$N rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/scale_n$N.json')); print('N', d['n_gpus'], 'ms', d['ms_per_step'], 'Mrays/s', d['value'], 'e2e', d['e2e']['value'], 'clocks', d['clocks'])"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 bench.py --gpus 8 --workload cfg5 --steps 5 --warmup 3 > gpurun_out/scale_cfg5_n8.json 2> gpurun_out/scale_cfg5_n8.err; echo "cfg5 rc=$?"; cut -c1-400 gpurun_out/scale_cfg5_n8.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29651 tools/stress_test.py --mode cem --iters 8 --rollouts 2048 > gpurun_out/stress_cem_n8.log 2>&1; echo "cem rc=$?"; tail -1 gpurun_out/stress_cem_n8.log | cut -c1-600
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29652 tools/stress_test.py --mode mc --rollouts 4096 > gpurun_out/stress_mc_n8.log 2>&1; echo "mc rc=$?"; tail -1 gpurun_out/stress_mc_n8.log | cut -c1-600
