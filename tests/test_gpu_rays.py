"""GPU: ray generation from poses -- ngp_get_rays against its numpy restatement (bit-exact) and the
reference's own get_rays output (tests/golden/pyref_cfg1.npz, <= 2e-7), and the pose-in / image-out C-ABI call."""
import os

import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import synthetic as syn
from nerfsafetyvalidation_b200.nerf.utils import get_rays
from oracle import run_sampler as rs

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "pyref_cfg1.npz"))


def test_get_rays_kernel_is_bit_exact_with_the_restatement_and_matches_the_reference(cuda):
    fx, fy, cx, cy, H, W = G["gr_intr"]
    H, W = int(H), int(W)
    r = get_rays(torch.from_numpy(G["gr_poses"]).to(cuda), (fx, fy, cx, cy), H, W)
    o, d = rs.get_rays(G["gr_poses"], fx, fy, cx, cy, H, W)
    assert np.array_equal(r["rays_o"].cpu().numpy(), o) and np.array_equal(r["rays_d"].cpu().numpy(), d)
    assert np.array_equal(r["rays_o"].cpu().numpy(), G["gr_rays_o"])
    assert np.abs(r["rays_d"].cpu().numpy() - G["gr_rays_d"]).max() <= 2e-7


def test_get_rays_kernel_full_frame_and_ragged_sizes(cuda):
    poses = np.stack([syn.orbit_pose(a) for a in (0.0, 37.0, 123.0, 250.0, 333.0)]).astype(np.float32)
    for H, W, f in ((800, 800, 1111.11), (7, 13, 9.5), (1, 1, 1.0), (200, 200, 277.78)):
        r = get_rays(torch.from_numpy(poses).to(cuda), (f, f, W / 2, H / 2), H, W)
        o, d = rs.get_rays(poses, f, f, W / 2, H / 2, H, W)
        assert np.array_equal(r["rays_d"].cpu().numpy(), d) and np.array_equal(r["rays_o"].cpu().numpy(), o)
    empty = get_rays(torch.zeros(0, 4, 4, device=cuda), (1.0, 1.0, 0.0, 0.0), 4, 4)
    assert empty["rays_o"].shape == (0, 16, 3)


def test_get_rays_rejects_cpu_poses_and_training_subsampling(cuda):
    with pytest.raises(RuntimeError):
        get_rays(torch.eye(4)[None], (1.0, 1.0, 0.0, 0.0), 4, 4)
    with pytest.raises(NotImplementedError):
        get_rays(torch.eye(4, device=cuda)[None], (1.0, 1.0, 0.0, 0.0), 4, 4, N=16)


def test_render_poses_host_equals_render_of_the_generated_rays(cuda, scene):
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant="B")
    model = syn.build_model(w, bitfield, grid, device=cuda)
    poses = np.stack([syn.orbit_pose(37.0), syn.orbit_pose(170.0, 1.3, 0.5), syn.orbit_pose(290.0)]).astype(np.float32)
    H = W = 40
    intr = (55.0, 55.0, W / 2, H / 2)
    dev_out = model.render_poses(torch.from_numpy(poses).to(cuda), intr, H, W, bg_color=1, perturb=False)
    img, dep = model.render_poses_host(torch.from_numpy(poses), intr, H, W)
    a, b = dev_out["image"].reshape(-1, 3).cpu().numpy(), img.numpy()
    assert np.array_equal(a, b)                                         # same kernels, same rays, same schedule
    da, db = dev_out["depth"].reshape(-1).cpu().numpy(), dep.numpy()
    assert np.array_equal(np.isnan(da), np.isnan(db)) and np.array_equal(da[~np.isnan(da)], db[~np.isnan(db)])
    # and against the oracle's renderer on the restated rays
    from oracle import oracle as orc
    o, d = rs.get_rays(poses, *intr, H, W)
    lib_scales = torch.empty(16, dtype=torch.float32, device=cuda)
    from nerfsafetyvalidation_b200 import _cabi
    _cabi.check(_cabi.load().ngp_grid_level_scales(16, w["S"], 16, _cabi.ptr(lib_scales), _cabi.stream()))
    c = orc.render(o.reshape(-1, 3), d.reshape(-1, 3), np.array([-1, -1, -1, 1, 1, 1], np.float32), bitfield, w["table"], w["offsets"], w["S"], 16,
                   w["w_sigma"], w["w_color"], density_scale=w["density_scale"], scales=lib_scales.cpu().numpy())
    assert np.abs(b - c["image"]).max() <= 1e-3                          # north-star tolerance for RGB
    with pytest.raises(RuntimeError):
        model.render_poses_host(torch.from_numpy(poses).to(cuda), intr, H, W)
