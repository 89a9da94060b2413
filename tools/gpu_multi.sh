# multi-GPU check: bench.py under torchrun at N = $NGPU (cfg2 and the cfg3 rollout batch), both arms
N=${NGPU:-2}
for args in "" "--workload rollout --variant B"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 8 --warmup 3 $args > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$? lines=$(wc -l < gpurun_out/bench_n$N.json)"
  python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print('N', d['n_gpus'], d['config']['workload'][:20], '| ms', round(d['ms_per_step'],3), 'Mrays/s', round(d['value'],2), 'e2e', round(d['e2e']['value'],2), 'pose', round(d['e2e_from_pose']['value'],2))"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2>/dev/null; echo "ref rc=$? lines=$(wc -l < gpurun_out/bench_ref_n$N.json)"; cut -c1-160 gpurun_out/bench_ref_n$N.json
