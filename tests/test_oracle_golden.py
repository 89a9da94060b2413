"""Pins the CPU oracle (oracle/ngp_oracle.c) against golden vectors produced by the reference's own CUDA extensions
rebuilt for sm_100a on a B200 (tests/golden/make_golden.py).  Runs without a GPU.

Bars: integer / index / bit outputs exact; march outputs (xyzs, dirs, deltas) bitwise; composite accumulators within
1e-6 (the reference uses MUFU ex2.approx, the CPU port exp2f); encoders / MLP within the stated tolerances.
"""
import os

import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import synthetic as syn
from oracle import oracle as orc
from tests.golden.make_golden import GOLDEN_FRAME, GRID_LOG2T, N_GRID, N_MLP, N_RAYS, N_SH
from tests.inputs import grid_inputs, make_rays, unit_dirs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    p = os.path.join(GOLD, name)
    if not os.path.exists(p):
        pytest.skip(f"{name} not generated yet (python tests/golden/make_golden.py on a GPU box)")
    return np.load(p)


def bitwise(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_setup_kernels():
    z = load("raymarch_setup.npz")
    o, d = make_rays(N_RAYS, 0)
    o, d = o.numpy(), d.numpy()
    for tag, b, mn in (("b1", 1.0, 0.2), ("b2", 2.0, 0.05)):
        n, f = orc.near_far_from_aabb(o, d, np.array([-b] * 3 + [b] * 3, np.float32), mn)
        assert bitwise(n, z["nears_" + tag]) and bitwise(f, z["fars_" + tag])
    assert np.array_equal(orc.morton3D(z["morton_coords"]), z["morton_idx"])
    assert np.array_equal(orc.morton3D_invert(z["invert_idx"]), z["invert_coords"])
    g = torch.Generator().manual_seed(21)
    torch.randint(0, 1024, (N_RAYS, 3), generator=g, dtype=torch.int32)
    torch.randint(0, 2 ** 30, (N_RAYS,), generator=g, dtype=torch.int32)
    grid = torch.rand(1, 65536, generator=g)
    grid[0, ::5] = 0.5
    assert np.array_equal(orc.packbits(grid.numpy(), 0.5), z["packbits_0p5"])


@pytest.mark.parametrize("variant", ["A", "B"])
def test_march_and_composite_snapshots(variant):
    z = load(f"render_{variant}.npz")
    bitfield, _ = syn.stonehenge_bitfield()
    o, d = syn.pinhole_rays(syn.orbit_pose(GOLDEN_FRAME["azimuth"]), GOLDEN_FRAME["H"], GOLDEN_FRAME["W"])
    nears, fars = orc.near_far_from_aabb(o, d, np.array([-1, -1, -1, 1, 1, 1], np.float32), 0.2)
    assert bitwise(nears, z["nears"]) and bitwise(fars, z["fars"])
    for i in z["picks"]:
        n_step = int(z[f"it{i}_n_step"])
        alive = z[f"it{i}_rays_alive"]
        n_alive = alive.shape[0]
        n = n_alive * n_step
        x, dd, dl = orc.march_rays(n_alive, n_step, alive, z[f"it{i}_rays_t"], o, d, 1.0, bitfield, 1, 128, nears, fars, 128, 0.0, 1024)
        assert x.shape[0] == int(z[f"it{i}_M"])
        assert bitwise(x[:n], z[f"it{i}_xyzs"]) and bitwise(dd[:n], z[f"it{i}_dirs"]) and bitwise(dl[:n], z[f"it{i}_deltas"])
        assert not x[n:].any() and not dl[n:].any()
        a = alive.copy()
        t, ws, dp, im = (z[f"it{i}_{k}"].copy() for k in ("rays_t", "weights_sum", "depth", "image"))
        orc.composite_rays(n_alive, n_step, a, t, z[f"it{i}_sigmas"], z[f"it{i}_rgbs"], z[f"it{i}_deltas"], ws, dp, im)
        assert np.array_equal(a, z[f"it{i}_alive_after"]), f"iteration {i}: alive flags"
        assert bitwise(t, z[f"it{i}_rays_t_after"])
        for got, k in ((ws, "weights_sum_after"), (dp, "depth_after"), (im, "image_after")):
            assert np.abs(got - z[f"it{i}_{k}"]).max() <= 1e-6
        assert np.array_equal(orc.compact_alive(a), a[a >= 0])


def test_march_random_cascade2():
    z = load("march_random.npz")
    o, d = make_rays(N_RAYS, 0)
    o2, d = (o * 2.0 / 2).numpy(), d.numpy()
    g = torch.Generator().manual_seed(11)
    bf = torch.randint(0, 256, (2 * 128 ** 3 // 8,), generator=g, dtype=torch.uint8).numpy()
    alive = torch.randperm(N_RAYS, generator=g)[:3000].int().numpy()
    assert np.array_equal(alive, z["alive"])
    nears, fars = orc.near_far_from_aabb(o2, d, np.array([-2] * 3 + [2] * 3, np.float32), 0.05)
    assert bitwise(nears, z["nears"]) and bitwise(fars, z["fars"])
    x, dd, dl = orc.march_rays(3000, 4, alive, z["rays_t"], o2, d, 2.0, bf, 2, 128, nears, fars, 128, 1.0 / 128, 1024)
    assert bitwise(x, z["xyzs"]) and bitwise(dd, z["dirs"]) and bitwise(dl, z["deltas"])
    assert (dl[:, 0] != 0).sum() > 1000


def test_encoders_and_mlp():
    z = load("encoders_mlp.npz")
    w = syn.ngp_weights(0, log2_hashmap_size=GRID_LOG2T, variant="B")
    x = grid_inputs(N_GRID, 5).numpy()
    y16, _ = orc.grid_encode(x, w["table"], w["offsets"], w["S"], 16, scales=z["scales"], round_mode=1)
    assert np.array_equal(y16.view(np.uint16), z["grid_f16"].view(np.uint16)), "fp16 table: reference rounding reproduced bit for bit"
    y32, _ = orc.grid_encode(x, w["table_f32"], w["offsets"], w["S"], 16, scales=z["scales"])
    assert bitwise(y32, z["grid_f32"])
    # the host-side exp2f differs from the device's ex2.approx by at most 1 ulp at these levels
    assert np.abs(orc.grid_level_scales(16, w["S"], 16) / z["scales"] - 1).max() < 3e-7
    dirs = unit_dirs(N_SH, 2)
    dirs[N_SH // 2:] *= 1.7
    assert np.allclose(orc.sh_encode(dirs.numpy(), 4), z["sh4"], atol=1e-5, rtol=1e-6)
    w19 = syn.ngp_weights(0)
    g = torch.Generator().manual_seed(34)
    xin = (torch.randn(N_MLP, 32, generator=g) * 0.5).half().numpy()
    for key, wk, nl in (("mlp_sigma", "w_sigma", 2), ("mlp_color", "w_color", 3)):
        got = orc.ffmlp(xin, w19[wk], 32, 16, 64, nl).astype(np.float32)
        ref = z[key].astype(np.float32)   # the reference accumulates in fp16; ours (and the port) in fp32
        assert np.abs(got - ref).max() <= 4e-3 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("variant", ["A", "B"])
def test_render_end_to_end(variant):
    z = load(f"render_{variant}.npz")
    w = syn.ngp_weights(0, variant=variant)
    bitfield, _ = syn.stonehenge_bitfield()
    o, d = syn.pinhole_rays(syn.orbit_pose(GOLDEN_FRAME["azimuth"]), GOLDEN_FRAME["H"], GOLDEN_FRAME["W"])
    sc = load("encoders_mlp.npz")["scales"]
    c = orc.render(o, d, np.array([-1, -1, -1, 1, 1, 1], np.float32), bitfield, w["table"], w["offsets"], w["S"], 16, w["w_sigma"], w["w_color"],
                   density_scale=w["density_scale"], scales=sc)
    assert np.abs(c["image"] - z["image"]).max() <= 1e-3
    ok = ~np.isnan(z["depth"])
    assert np.array_equal(np.isnan(c["depth"]), ~ok) and np.abs(c["depth"][ok] - z["depth"][ok]).max() <= 1e-3
    n_diff = int((c["samples_per_ray"] != z["samples_per_ray"]).sum())
    if variant == "A":
        assert n_diff == 0 and c["iterations"] == int(z["iterations"]) and c["samples"] == int(z["samples"])
    else:
        assert n_diff <= 8
