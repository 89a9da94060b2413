# BASELINE.json configs[2..4] on N GPUs of one box:  tools/gpu_scale_stress.sh N   (run under `gpurun --gpus N`)
#   cfg3  Monte Carlo: 4096 rollouts x 20 frames of 200x200        cfg4  CEM: 8 iterations x 2048 rollouts, elite statistics over NCCL
#   cfg5  ONE 1920x1080 frame, 2^22 table, dense occupancy, split by row blocks across the ranks (strong scaling)
N=${1:-1}
RUN="python"; [ "$N" -gt 1 ] && RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
set -x
timeout 600 $RUN tools/stress_test.py --mode mc --rollouts 4096 --steps 20 --res 200 > gpurun_out/r02_scale_cfg3_mc_n$N.json 2> gpurun_out/scale_mc_n$N.err; tail -c 300 gpurun_out/r02_scale_cfg3_mc_n$N.json
timeout 900 $RUN tools/stress_test.py --mode cem --iters 8 --rollouts 2048 --steps 20 --res 200 > gpurun_out/r02_scale_cfg4_cem_n$N.json 2> gpurun_out/scale_cem_n$N.err; tail -c 300 gpurun_out/r02_scale_cfg4_cem_n$N.json
SPLIT=; [ "$N" -gt 1 ] && SPLIT=--split-rows
timeout 600 $RUN bench.py --gpus $N --workload cfg5 $SPLIT --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/r02_scale_cfg5_rows_n$N.json 2> gpurun_out/scale_cfg5_n$N.err; python -c "
import json; d=json.load(open('gpurun_out/r02_scale_cfg5_rows_n$N.json')); print('CFG5 n=$N ms', d['ms_per_step'], 'Mrays/s', d['value'])"
timeout 300 $RUN bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/r02_scale_cfg2_n$N.json 2> gpurun_out/scale_cfg2_n$N.err; python -c "
import json; d=json.load(open('gpurun_out/r02_scale_cfg2_n$N.json')); print('CFG2 n=$N ms', d['ms_per_step'], 'Mrays/s', d['value'], 'e2e', d['e2e']['value'])"
