"""CPU: oracle/run_sampler.py (the restatement of the reference's pure-PyTorch `run()` path, BASELINE configs[0], and of
nerf/utils.py:get_rays) against tests/golden/pyref_cfg1.npz, which tests/golden/make_golden_pyref.py produced by importing and
running the reference's OWN Python on CPU."""
import os

import numpy as np
import torch

from oracle import run_sampler as rs

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "pyref_cfg1.npz"))


def test_seeded_init_reproduces_the_reference_weights():
    net = rs.VanillaNeRF(seed=0)
    assert net.n_params() == int(G["n_params"]) == 11264
    for k, w in enumerate(net.sigma):
        assert np.array_equal(w.numpy(), G[f"sigma_w{k}"])
    for k, w in enumerate(net.color):
        assert np.array_equal(w.numpy(), G[f"color_w{k}"])


def test_get_rays_restatement_matches_the_reference():
    fx, fy, cx, cy, H, W = G["gr_intr"]
    o, d = rs.get_rays(G["gr_poses"], fx, fy, cx, cy, int(H), int(W))
    assert np.array_equal(o, G["gr_rays_o"])                       # origins are copies of the pose translation
    assert np.abs(d - G["gr_rays_d"]).max() <= 2e-7                # torch.norm + matmul vs our fixed operation order: <= 1-2 ulp
    assert np.abs(np.linalg.norm(d, axis=-1) - 1).max() <= 3e-7


def test_run_restatement_matches_the_reference_frame():
    Hc, Wc = (int(v) for v in G["cfg1_hw"])
    fl = float(G["cfg1_focal"])
    o, d = rs.get_rays(G["cfg1_pose"][None], fl, fl, Wc / 2, Hc / 2, Hc, Wc)
    net = rs.VanillaNeRF(seed=0)
    out = rs.render(net, torch.from_numpy(o[0]), torch.from_numpy(d[0]), max_ray_batch=300, num_steps=64, upsample_steps=64, bg_color=1.0)
    # tolerance 2e-5 abs: the only difference is the 1-ulp ray directions feeding sin / cos of up to 32 x
    assert np.abs(out["image"].numpy() - G["cfg1_image"]).max() <= 2e-5
    assert np.abs(out["depth"].numpy() - G["cfg1_depth"]).max() <= 2e-5
    assert np.abs(out["aggregated_density"].numpy() - G["cfg1_aggregated_density"]).max() <= 2e-4 * max(1.0, float(np.abs(G["cfg1_aggregated_density"]).max()))
    assert 0.01 < float(out["weights_sum"].mean()) < 1.0           # the frame is not trivially empty / saturated


def test_chunking_does_not_change_the_result():
    o, d = rs.get_rays(G["cfg1_pose"][None], 40.0, 40.0, 8.0, 8.0, 16, 16)
    net = rs.VanillaNeRF(seed=0)
    a = rs.render(net, torch.from_numpy(o[0]), torch.from_numpy(d[0]), max_ray_batch=4096)
    b = rs.render(net, torch.from_numpy(o[0]), torch.from_numpy(d[0]), max_ray_batch=50)
    assert np.abs(a["image"].numpy() - b["image"].numpy()).max() <= 1e-6


def test_bench_poses_reproduce_the_bench_rays():
    """bench.py hands the same frames to the ray API and to the pose API (e2e_from_pose): the poses must regenerate the rays."""
    import bench
    wl = dict(bench.WORKLOADS["rollout"], H=20, W=24, focal=30.0)
    o, d, poses = bench.frame_rays(wl, 0, 5, with_poses=True)
    o2, d2 = rs.get_rays(poses, wl["focal"], wl["focal"], wl["W"] / 2, wl["H"] / 2, wl["H"], wl["W"])
    assert poses.shape == (5, 4, 4) and np.array_equal(o2.reshape(-1, 3), o) and np.abs(d2.reshape(-1, 3) - d).max() <= 2e-7
    wl = dict(bench.WORKLOADS["cfg2"], H=16, W=16, focal=22.0)
    o, d, poses = bench.frame_rays(wl, 3, 1, with_poses=True)
    o2, d2 = rs.get_rays(poses, 22.0, 22.0, 8.0, 8.0, 16, 16)
    assert np.array_equal(o2.reshape(-1, 3), o) and np.abs(d2.reshape(-1, 3) - d).max() <= 2e-7


def test_cfg1_full_frame_reproduces_the_reference_probe():
    """BASELINE configs[0] at its full size (100x100, 64+64 samples, seed 0): SURVEY.md Appendix C records the mean pixel the
    reference's own run() produced for it, 0.59923."""
    import math
    H = W = 100
    fl = 0.5 * W / math.tan(0.5 * 0.6911112)
    pose = np.eye(4, dtype=np.float32)
    pose[2, 3] = -2.5
    o, d = rs.get_rays(pose[None], fl, fl, W / 2, H / 2, H, W)
    out = rs.render(rs.VanillaNeRF(seed=0), torch.from_numpy(o[0]), torch.from_numpy(d[0]), max_ray_batch=4096, num_steps=64,
                    upsample_steps=64, bg_color=1.0)
    assert abs(float(out["image"].mean()) - 0.59923) <= 1e-5
