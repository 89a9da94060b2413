"""Generates tests/golden/pyref_cfg1.npz by IMPORTING the reference's own Python (CPU, no GPU needed):

    python tests/golden/make_golden_pyref.py            (needs /root/reference; writes tests/golden/pyref_cfg1.npz)

What is imported untouched from the reference tree: nerf/renderer.py (NeRFRenderer.run / render, sample_pdf), nerf/network.py
(the vanilla NeRFNetwork), encoding.py (FreqEncoder, get_encoder), activation.py (trunc_exp), and the source text of
nerf/utils.py:get_rays / custom_meshgrid (nerf/utils.py itself imports tensorboardX, mcubes, lpips, ... which are absent).
Stubs (SURVEY.md section 8c recipe): `trimesh` (only used by plot_pointcloud), `raymarching` with near_far_from_aabb answered by
the C oracle (itself pinned bitwise against the reference's CUDA kernel, tests/golden/raymarch_setup.npz), `freqencoder` mapped to
encoding.FreqEncoder(max_freq_log2=degree-1, N_freqs=degree) -- the equivalence the reference states at encoding.py:55.

The fixture pins oracle/run_sampler.py (BASELINE configs[0], the CPU baseline path) and the get_rays restatement / kernel.
"""
import ast
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("NGP_REFERENCE_ROOT", "/root/reference")
sys.path.insert(0, ROOT)

from oracle import oracle as orc  # noqa: E402


def import_reference():
    sys.modules["trimesh"] = types.ModuleType("trimesh")
    rm = types.ModuleType("raymarching")

    def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
        n, f = orc.near_far_from_aabb(rays_o.detach().numpy(), rays_d.detach().numpy(), aabb.numpy(), float(min_near))
        return torch.from_numpy(n), torch.from_numpy(f)

    rm.near_far_from_aabb = near_far_from_aabb
    sys.modules["raymarching"] = rm
    sys.path.insert(0, REF)
    import encoding  # the reference's encoding.py
    fe = types.ModuleType("freqencoder")
    fe.FreqEncoder = lambda input_dim=3, degree=4: encoding.FreqEncoder(input_dim=input_dim, max_freq_log2=degree - 1, N_freqs=degree)
    sys.modules["freqencoder"] = fe
    # nerf package: real renderer.py / network.py, utils reduced to the two functions on the path (source text from the reference)
    src = open(os.path.join(REF, "nerf", "utils.py")).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("custom_meshgrid", "get_rays")]
    utils = types.ModuleType("nerf.utils")
    utils.torch = torch
    utils.np = np
    from packaging import version as pver
    utils.pver = pver
    exec(compile(ast.Module(body=keep, type_ignores=[]), os.path.join(REF, "nerf", "utils.py"), "exec"), utils.__dict__)
    pkg = types.ModuleType("nerf")
    pkg.__path__ = [os.path.join(REF, "nerf")]
    sys.modules["nerf"] = pkg
    sys.modules["nerf.utils"] = utils
    from nerf.network import NeRFNetwork
    return NeRFNetwork, utils.get_rays


def main():
    NeRFNetwork, get_rays = import_reference()
    from nerfsafetyvalidation_b200 import synthetic as syn
    torch.manual_seed(0)
    model = NeRFNetwork(encoding="frequency", encoding_dir="frequency", bound=1, cuda_ray=False, density_scale=1, min_near=0.2,
                        density_thresh=10, bg_radius=-1).eval()
    n_params = sum(p.numel() for p in model.parameters())
    assert n_params == 11264, n_params
    out = {"n_params": np.int64(n_params)}
    for k, lin in enumerate(model.sigma_net):
        out[f"sigma_w{k}"] = lin.weight.detach().numpy()
    for k, lin in enumerate(model.color_net):
        out[f"color_w{k}"] = lin.weight.detach().numpy()
    # --- get_rays: three poses (cfg2 orbit + two tilted ones), 12 x 16 pixels
    poses = np.stack([syn.orbit_pose(37.0), syn.orbit_pose(200.0, 1.2, 0.6), syn.lookat_pose((0.3, 0.2, -2.5), up=(0.1, 1.0, 0.0))]).astype(np.float32)
    H, W, focal = 12, 16, 20.0
    r = get_rays(torch.from_numpy(poses), (focal, focal * 1.1, W / 2, H / 2 + 0.25), H, W, -1)
    out.update(gr_poses=poses, gr_intr=np.array([focal, focal * 1.1, W / 2, H / 2 + 0.25, H, W], np.float64),
               gr_rays_o=r["rays_o"].contiguous().numpy(), gr_rays_d=r["rays_d"].contiguous().numpy())
    # --- cfg1-shaped frame (32 x 32 instead of 100 x 100 so the fixture stays small), camera at (0,0,-2.5) looking +z
    Hc = Wc = 32
    fl = 0.5 * Wc / np.tan(0.5 * 0.6911112)
    pose = np.eye(4, dtype=np.float32)
    pose[2, 3] = -2.5
    rr = get_rays(torch.from_numpy(pose[None]), (fl, fl, Wc / 2, Hc / 2), Hc, Wc, -1)
    with torch.no_grad():
        res = model.render(rr["rays_o"].contiguous(), rr["rays_d"].contiguous(), staged=True, bg_color=1.0, perturb=False, num_steps=64,
                           upsample_steps=64, max_ray_batch=300)
    out.update(cfg1_pose=pose, cfg1_hw=np.array([Hc, Wc], np.int64), cfg1_focal=np.float64(fl), cfg1_image=res["image"][0].numpy(),
               cfg1_depth=res["depth"][0].numpy(), cfg1_aggregated_density=res["aggregated_density"][0].numpy())
    # the model's weights again from the same seed: oracle/run_sampler.VanillaNeRF(seed=0) must reproduce them without the fixture
    dst = os.path.join(ROOT, "tests", "golden", "pyref_cfg1.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
