# what the driver runs at round end, in its order: GPU tests, smoke, reference arm, our arm (defaults)
set -x
timeout 900 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -3
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 5 --warmup 1 > gpurun_out/drv_ref.json 2> gpurun_out/drv_ref.err ) 2>&1 | grep real; wc -l gpurun_out/drv_ref.json; cut -c1-300 gpurun_out/drv_ref.json
( time timeout 900 python bench.py > gpurun_out/drv_ours.json 2> gpurun_out/drv_ours.err ) 2>&1 | grep real; wc -l gpurun_out/drv_ours.json; cut -c1-200 gpurun_out/drv_ours.json; tail -3 gpurun_out/drv_ours.err
