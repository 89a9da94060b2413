"""CPU: the bench.py contract the driver relies on -- the reference arm prints exactly ONE JSON line with the agreed keys, and
the product arm refuses to run without a CUDA device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*args):
    env = dict(os.environ, OMP_NUM_THREADS="1")   # what torchrun exports; the arm must override it
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)


def test_reference_arm_prints_one_json_line():
    r = run_bench("--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-rays", "1500")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["metric"].startswith("rendered Mrays/s") and d["value"] > 0 and d["config"]["workload"].startswith("cfg2")
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == (os.cpu_count() or 1) and cb["value"] == d["value"] and "rays" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    run = d["cpu_reference_run"]          # BASELINE configs[0]: the reference's pure-PyTorch run() on the host cores, beside it
    assert run["value"] > 0 and run["cores"] == (os.cpu_count() or 1) and "cfg1" in run["what"]


def test_reference_arm_does_nothing_on_other_ranks():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"], capture_output=True, text=True,
                       env=env, cwd=ROOT, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


@pytest.mark.skipif(torch.cuda.is_available(), reason="this check is about a machine WITHOUT a GPU")
def test_product_arm_has_no_cpu_fallback():
    r = run_bench("--steps", "1", "--warmup", "1")
    assert r.returncode != 0 and r.stdout.strip() == "" and "no CUDA device" in r.stderr
