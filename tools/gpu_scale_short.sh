# scaling record, bench only: N = 1, 2, 4, 8 on one box (cfg2)
for N in 1 2 4 8; do
  if [ $N = 1 ]; then timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; fi
  echo "N=$N rc=$? lines=$(wc -l < gpurun_out/scale_n$N.json)"; python -c "
import json; d=json.load(open('gpurun_out/scale_n$N.json')); print('N', d['n_gpus'], 'ms', d['ms_per_step'], 'Mrays/s', d['value'], 'e2e', d['e2e']['value'], 'clocks', d['clocks'])"
done
