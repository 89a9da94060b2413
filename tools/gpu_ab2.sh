bash tools/gpu_ab.sh base3 cur base3 cur 2>&1 | grep "^AB"
timeout 200 python tools/parity_margin.py 2>&1 | tail -20
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
