"""GPU: parity on the BASELINE.json configurations the bench numbers are quoted on, each against the reference's own extensions
rebuilt for sm_100a (tests/refext.RefModel.run_cuda) with IDENTICAL batching (SURVEY.md Appendix C, row "render e2e"):

  cfg3 batch   64 rollout frames of 200x200 (2.56 M rays) in ONE render call, variant-B weights (rays terminate by T < 1e-4):
               the shape every rollout / Monte Carlo / CEM number is quoted on
  cfg2 B       the 800x800 frame with variant-B weights (epoch speculation with transmittance deaths at full size)
  cfg5         the whole 1920x1080 frame, 2^22 table (151 MB fp16: the fine levels leave L2), all-ones occupancy, max_steps 1024
  NCCL         rollouts.cem_update on two GPUs (one process per GPU, `nccl`): the same proposal on both ranks, equal to the
               single-process update (skipped with fewer than two devices)

Integer outputs (per-ray sample counts, iteration count) are exact where nothing terminates by transmittance (variant A) and
exact up to the enumerated rays that sit on the T < 1e-4 threshold otherwise (fp32- vs fp16-accumulated sigma); images, depth
and weights_sum within the north-star tolerance 1e-3 abs.
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import rollouts as ro, synthetic as syn
from tests import refext

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not refext.available("_ffmlp"), reason="oracle/_ref/*.so not built (python oracle/build_ref.py)")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-3   # north_star: RGB / depth within 1e-3 abs


def render(model, o, d, prefix=None):
    with torch.no_grad():
        o = o[None] if prefix is None else o.view(*prefix, 3)
        d = d[None] if prefix is None else d.view(*prefix, 3)
        return model.render(o, d, bg_color=1, perturb=False, return_stats=True)


def compare(out, ref, n_rays, exact_counts, what):
    """out: our result dict (flattened below), ref: RefModel.run_cuda's.  Returns the number of rays whose sample count differs."""
    img, dep, ws = out["image"].reshape(-1, 3), out["depth"].reshape(-1), out["weights_sum"].reshape(-1)
    assert img.shape[0] == n_rays
    stats = out["stats"].cpu().tolist()
    assert stats[2] >= 0, f"{what}: the renderer's internal consistency check tripped"
    assert torch.equal(torch.isnan(dep), torch.isnan(ref["depth"])), f"{what}: NaN-depth pattern of missed rays (renderer.py:376)"
    hit = ~torch.isnan(dep)
    e_img = (img - ref["image"]).abs().max().item()
    e_dep = (dep - ref["depth"])[hit].abs().max().item()
    e_ws = (ws - ref["weights_sum"]).abs().max().item()
    assert e_img <= TOL and e_dep <= TOL and e_ws <= TOL, f"{what}: image {e_img:.2e} depth {e_dep:.2e} weights_sum {e_ws:.2e}"
    n_diff = int((out["samples_per_ray"] != ref["samples_per_ray"]).sum())
    if exact_counts:
        assert n_diff == 0 and stats[0] == ref["iterations"] and stats[1] == ref["samples"], \
            f"{what}: {n_diff} rays differ, iterations {stats[0]} vs {ref['iterations']}, samples {stats[1]} vs {ref['samples']}"
    else:
        # threshold rays: T crosses 1e-4 one sample earlier / later with fp32- vs fp16-accumulated sigma; enumerated, bounded
        assert n_diff <= max(8, n_rays // 500), f"{what}: {n_diff} of {n_rays} rays differ in sample count"
        assert abs(stats[0] - ref["iterations"]) <= 2, f"{what}: iterations {stats[0]} vs {ref['iterations']}"
        assert abs(stats[1] - ref["samples"]) <= max(64, ref["samples"] // 2000)
    return n_diff


def rollout_batch(n_frames=64, steps=20, H=200, W=200, focal=277.78):
    """bench.py --workload rollout, rank 0: the first n_frames frames of rollouts 0, 1, 2, ... (seed = rollout id)."""
    ids = ro.shard_ids(10 ** 6, 0, 1)[: -(-n_frames // steps)]
    o, d, _, _ = ro.rollout_rays(ids, steps, H, W, focal)
    return np.ascontiguousarray(o[:n_frames]), np.ascontiguousarray(d[:n_frames])


@needs_ref
def test_cfg3_rollout_batch_against_the_reference_extensions(cuda, scene):
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant="B")
    o, d = rollout_batch()
    F, P = o.shape[:2]
    o, d = torch.from_numpy(o).to(cuda), torch.from_numpy(d).to(cuda)
    ref = refext.RefModel(w, bitfield, device=cuda).run_cuda(o.view(-1, 3), d.view(-1, 3))   # renderer.py:265-269 flattens [B,N,3]
    out = render(syn.build_model(w, bitfield, grid, device=cuda), o, d, prefix=(F, P))
    assert out["image"].shape == (F, P, 3) and out["depth"].shape == (F, P)
    compare(out, ref, F * P, exact_counts=False, what="cfg3 batch (64 x 200x200, variant B)")


@needs_ref
def test_cfg2_variant_b_full_frame_against_the_reference_extensions(cuda, scene):
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant="B")
    o, d = syn.pinhole_rays(syn.orbit_pose(), 800, 800, focal=1111.11)
    o, d = torch.from_numpy(o).to(cuda), torch.from_numpy(d).to(cuda)
    ref = refext.RefModel(w, bitfield, device=cuda).run_cuda(o, d)
    out = render(syn.build_model(w, bitfield, grid, device=cuda), o, d)
    compare(out, ref, 640000, exact_counts=False, what="cfg2 variant B (800x800)")
    assert (out["weights_sum"] > 0.9999).sum().item() > 1000, "variant B is supposed to terminate rays by transmittance"


@needs_ref
def test_cfg5_full_frame_against_the_reference_extensions(cuda, scene):
    _, grid = scene
    bitfield = np.full(128 ** 3 // 8, 255, np.uint8)
    w = syn.ngp_weights(0, log2_hashmap_size=22, variant="A")
    o, d = syn.pinhole_rays(syn.orbit_pose(), 1080, 1920, focal=1500.0)
    o, d = torch.from_numpy(o).to(cuda), torch.from_numpy(d).to(cuda)
    ref = refext.RefModel(w, bitfield, device=cuda).run_cuda(o, d)
    out = render(syn.build_model(w, bitfield, grid, device=cuda), o, d)
    compare(out, ref, 1920 * 1080, exact_counts=True, what="cfg5 (1920x1080, 2^22 table, dense occupancy)")
    assert ref["samples"] > 10 ** 9   # ~1.2 G samples per frame


# ---------------------------------------------------------------------------------------------- NCCL, two GPUs
_WORKER = r"""
import json, os, sys
sys.path.insert(0, {root!r})
import numpy as np, torch, torch.distributed as dist
from nerfsafetyvalidation_b200 import rollouts as ro
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
n, steps, m_elite = 64, 5, 8
p_mean = np.zeros((steps, 12)); p_std = np.broadcast_to(np.asarray(ro.ENV["noise_std"]), (steps, 12)).copy()
ids = ro.shard_ids(n, rank, world)
noise = ro.sample_noise(ids, steps)
risks = torch.from_numpy(np.sin(ids * 0.37).astype(np.float32)).to(dev)          # a deterministic stand-in for rendered risks
mean, var, elite = ro.cem_update(risks, torch.from_numpy(noise), ids, n, m_elite, p_mean, p_std, p_mean, p_std * 1.5, rank, world)
both = [torch.empty_like(mean) for _ in range(world)]
dist.all_gather(both, mean)
same = all(torch.equal(both[0], b) for b in both)
if rank == 0:
    print("RESULT " + json.dumps(dict(mean=mean.cpu().tolist(), var=var.cpu().tolist(), elite=elite, same=same, backend=dist.get_backend())))
dist.destroy_process_group()
"""


@pytest.mark.skipif(torch.cuda.is_available() and torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_cem_update_over_nccl_on_two_gpus(cuda, tmp_path):
    """CrossEntropyMethod.py:211-251 sharded: all-gather of risks + all-reduce(max) + all-reduce(sum) of the weighted moments over
    NCCL give both ranks the single-process update."""
    script = tmp_path / "cem_worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29731", str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads(next(ln for ln in r.stdout.splitlines() if ln.startswith("RESULT "))[7:])
    assert res["backend"] == "nccl" and res["same"]
    n, steps, m_elite = 64, 5, 8
    p_mean = np.zeros((steps, 12))
    p_std = np.broadcast_to(np.asarray(ro.ENV["noise_std"]), (steps, 12)).copy()
    ids = ro.shard_ids(n, 0, 1)
    risks = torch.from_numpy(np.sin(ids * 0.37).astype(np.float32)).to(cuda)
    mean, var, elite = ro.cem_update(risks, torch.from_numpy(ro.sample_noise(ids, steps)), ids, n, m_elite, p_mean, p_std, p_mean, p_std * 1.5)
    assert sorted(elite) == sorted(res["elite"])
    assert np.allclose(np.asarray(res["mean"]), mean.cpu().numpy(), rtol=0, atol=1e-12)
    assert np.allclose(np.asarray(res["var"]), var.cpu().numpy(), rtol=0, atol=1e-12)
