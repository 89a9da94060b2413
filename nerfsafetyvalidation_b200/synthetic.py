"""Deterministic synthetic inputs shaped like the reference's Stonehenge deployment.

numpy only, so the CPU oracle tests, the GPU parity tests and bench.py all see the same bytes:
  * `stonehenge_density_grid`  analytic occupancy in [-1,1]^3 sampled at the 128^3 cell centres, Morton order like
                               NeRFRenderer.density_grid (nerf/renderer.py:86-90, raymarching.cu:58-73)
  * `orbit_pose`, `pinhole_rays`  camera-to-world pose in the NGP frame + rays as nerf/utils.py:52-116 builds them
  * `ngp_weights`              hash-grid table / MLP weights with the reference initialisers' distributions
                               (grid.py:133-135, ffmlp.py:141-144), seeded numpy generators
  * `hashgrid_offsets`         per-level row offsets exactly as gridencoder/grid.py:112-127
"""
import math

import numpy as np


# ------------------------------------------------------------------ Morton (raymarching.cu:58-73)
def _expand_bits(v):
    v = v.astype(np.uint32)
    v = (v * np.uint32(0x00010001)) & np.uint32(0xFF0000FF)
    v = (v * np.uint32(0x00000101)) & np.uint32(0x0F00F00F)
    v = (v * np.uint32(0x00000011)) & np.uint32(0xC30C30C3)
    v = (v * np.uint32(0x00000005)) & np.uint32(0x49249249)
    return v


def morton3d(x, y, z):
    return _expand_bits(x) | (_expand_bits(y) << np.uint32(1)) | (_expand_bits(z) << np.uint32(2))


# ------------------------------------------------------------------ scene
def _box(p, lo, hi):
    return np.all((p >= np.asarray(lo, np.float32)) & (p <= np.asarray(hi, np.float32)), axis=-1)


def _rot_box(p, centre_xy, angle, half, z0, z1):
    """Axis box of half-extents `half` (radial, tangential) rotated by `angle` about z, spanning z0..z1 (world frame)."""
    c, s = math.cos(angle), math.sin(angle)
    dx, dy = p[..., 0] - centre_xy[0], p[..., 1] - centre_xy[1]
    u = c * dx + s * dy
    v = -s * dx + c * dy
    return (np.abs(u) <= half[0]) & (np.abs(v) <= half[1]) & (p[..., 2] >= z0) & (p[..., 2] <= z1)


def stonehenge_inside(p_world):
    """p_world [...,3] (x, y, z-up) -> bool: ground slab + outer ring of 30 pillars with lintels + inner horseshoe of
    5 trilithons."""
    p = np.asarray(p_world, np.float32)
    inside = _box(p, (-1, -1, -0.1), (1, 1, 0.0))
    n_outer, r_outer = 30, 0.62
    for k in range(n_outer):
        a = 2 * math.pi * k / n_outer
        inside |= _rot_box(p, (r_outer * math.cos(a), r_outer * math.sin(a)), a, (0.03, 0.04), 0.0, 0.45)
        a2 = a + math.pi / n_outer   # lintel spanning to the next pillar
        inside |= _rot_box(p, (r_outer * math.cos(a2), r_outer * math.sin(a2)), a2, (0.03, 0.07), 0.45, 0.51)
    r_inner = 0.32
    for k in range(5):
        a = math.radians(-120 + 60 * k)
        for da in (-0.16, 0.16):
            inside |= _rot_box(p, (r_inner * math.cos(a + da), r_inner * math.sin(a + da)), a + da, (0.04, 0.05), 0.0, 0.6)
        inside |= _rot_box(p, (r_inner * math.cos(a), r_inner * math.sin(a)), a, (0.04, 0.11), 0.6, 0.68)
    return inside


def stonehenge_density_grid(grid_size=128, cascade=1, value=1000.0):
    """float32 [cascade, grid_size^3] in Morton order; cascade c covers [-2^c, 2^c]^3 (clipped scene repeats nothing:
    the scene only lives in [-1,1]^3)."""
    H = grid_size
    idx = np.arange(H, dtype=np.uint32)
    ix, iy, iz = np.meshgrid(idx, idx, idx, indexing="ij")
    m = morton3d(ix.ravel(), iy.ravel(), iz.ravel()).astype(np.int64)
    grid = np.zeros((cascade, H ** 3), np.float32)
    for c in range(cascade):
        b = float(2 ** c)
        ctr = ((np.stack([ix, iy, iz], -1).reshape(-1, 3).astype(np.float32) * 2 + 1) / H - 1) * b   # NGP-frame cell centres
        # NGP frame (x,y,z)_ngp = (y,z,x)_world (nerf/provider.py:19-27)  =>  world = (z_ngp, x_ngp, y_ngp)
        world = np.stack([ctr[:, 2], ctr[:, 0], ctr[:, 1]], -1)
        grid[c, m] = np.where(stonehenge_inside(world), value, 0.0).astype(np.float32)
    return grid


def packbits_np(grid, thresh):
    """raymarching.cu:269-291 on the host (for building scenes without a GPU): LSB-first."""
    g = (np.asarray(grid, np.float32).reshape(-1, 8) > np.float32(thresh)).astype(np.uint8)
    return (g << np.arange(8, dtype=np.uint8)).sum(axis=1).astype(np.uint8)


def stonehenge_bitfield(grid_size=128, cascade=1, density_thresh=10.0):
    """density_bitfield exactly as renderer.py:533-538 derives it: thresh = min(mean(clamp(grid, 0)), density_thresh)."""
    grid = stonehenge_density_grid(grid_size, cascade)
    mean_density = float(np.mean(np.clip(grid, 0, None)))
    return packbits_np(grid, min(mean_density, density_thresh)), grid


# ------------------------------------------------------------------ cameras
def lookat_pose(eye, target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0)):
    """4x4 cam2world whose columns are (right, down, forward, eye) -- the convention nerf/utils.py:103-111 consumes
    (direction = (x, y, 1) with y growing down the image).  NGP frame: +y is up (world z)."""
    eye = np.asarray(eye, np.float64)
    f = np.asarray(target, np.float64) - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, np.asarray(up, np.float64))
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    pose = np.eye(4, dtype=np.float32)
    pose[:3, 0], pose[:3, 1], pose[:3, 2], pose[:3, 3] = r, d, f, eye
    return pose


def orbit_pose(azimuth_deg=37.0, radius=1.6, height=0.35):
    a = math.radians(azimuth_deg)
    # world (r cos a, r sin a, h) -> NGP (y, z, x)_world
    return lookat_pose((radius * math.sin(a), height, radius * math.cos(a)))


def pinhole_rays(pose, H, W, focal=None, fovx=0.6911112):
    """rays_o, rays_d float32 [H*W,3] as nerf/utils.py:69-71,103-111 (pixel centres +0.5, normalised directions)."""
    if focal is None:
        focal = 0.5 * W / math.tan(0.5 * fovx)
    cx, cy = W / 2, H / 2
    i, j = np.meshgrid(np.arange(W, dtype=np.float32) + np.float32(0.5), np.arange(H, dtype=np.float32) + np.float32(0.5))
    xs = ((i.reshape(-1) - np.float32(cx)) / np.float32(focal)).astype(np.float32)
    ys = ((j.reshape(-1) - np.float32(cy)) / np.float32(focal)).astype(np.float32)
    d = np.stack([xs, ys, np.ones_like(xs)], -1)
    d = d / np.linalg.norm(d, axis=-1, keepdims=True).astype(np.float32)
    R = np.asarray(pose, np.float32)[:3, :3]
    rays_d = (d @ R.T).astype(np.float32)
    rays_o = np.broadcast_to(np.asarray(pose, np.float32)[:3, 3], rays_d.shape).copy()
    return rays_o, np.ascontiguousarray(rays_d)


# ------------------------------------------------------------------ network
def hashgrid_offsets(num_levels=16, level_dim=2, base_resolution=16, log2_hashmap_size=19, desired_resolution=2048,
                     input_dim=3, align_corners=False):
    """(offsets int32 [L+1], per_level_scale) as gridencoder/grid.py:99,112-127."""
    per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
    max_params = 2 ** log2_hashmap_size
    offsets, offset = [], 0
    for i in range(num_levels):
        resolution = int(np.ceil(base_resolution * per_level_scale ** i))
        n = min(max_params, (resolution if align_corners else resolution + 1) ** input_dim)
        n = int(np.ceil(n / 8) * 8)
        offsets.append(offset)
        offset += n
    offsets.append(offset)
    return np.asarray(offsets, np.int32), float(per_level_scale)


def ngp_weights(seed=0, log2_hashmap_size=19, bound=1, variant="A", num_layers=2, num_layers_color=3):
    """dict(table fp16 [rows,2], offsets, per_level_scale, S, w_sigma fp16, w_color fp16, density_scale).
    variant A: U(+-1e-4) table (reference init), density_scale 1.  variant B: table x 1e4, density_scale 50, so sigma
    spans orders of magnitude and rays terminate by transmittance."""
    offsets, pls = hashgrid_offsets(log2_hashmap_size=log2_hashmap_size, desired_resolution=2048 * bound)
    rng = np.random.default_rng(seed)
    table = rng.uniform(-1e-4, 1e-4, size=(int(offsets[-1]), 2)).astype(np.float32)
    density_scale = 1.0
    if variant == "B":
        table *= np.float32(1e4)
        density_scale = 50.0
    std = math.sqrt(3 / 64)
    rng_w = np.random.default_rng(42 + seed)
    n_sigma = 64 * (32 + 64 * (num_layers - 1) + 16)
    n_color = 64 * (32 + 64 * (num_layers_color - 1) + 16)
    w_sigma = rng_w.uniform(-std, std, size=n_sigma).astype(np.float32)
    w_color = rng_w.uniform(-std, std, size=n_color).astype(np.float32)
    return dict(table=table.astype(np.float16), table_f32=table, offsets=offsets, per_level_scale=pls, S=float(np.log2(pls)),
                H=16, w_sigma=w_sigma.astype(np.float16), w_color=w_color.astype(np.float16), w_sigma_f32=w_sigma,
                w_color_f32=w_color, density_scale=density_scale, bound=float(bound), num_layers=num_layers,
                num_layers_color=num_layers_color)


def build_model(weights, bitfield=None, density_grid=None, device="cuda", min_near=0.2, density_thresh=10.0):
    """A `nerf.network_ff.NeRFNetwork` (ours) carrying `weights` and the scene, in eval mode on `device`."""
    import torch

    from .nerf.network_ff import NeRFNetwork
    log2T = int(round(math.log2(int(weights["offsets"][-1] - weights["offsets"][-2]))))
    model = NeRFNetwork(encoding="hashgrid", bound=weights["bound"], cuda_ray=True, density_scale=weights["density_scale"],
                        min_near=min_near, density_thresh=density_thresh, bg_radius=-1, log2_hashmap_size=log2T,
                        num_layers=weights["num_layers"], num_layers_color=weights["num_layers_color"])
    with torch.no_grad():
        model.encoder.embeddings.copy_(torch.from_numpy(weights["table_f32"]))
        model.sigma_net.weights.copy_(torch.from_numpy(weights["w_sigma_f32"]))
        model.color_net.weights.copy_(torch.from_numpy(weights["w_color_f32"]))
        if density_grid is not None:
            model.density_grid.copy_(torch.from_numpy(density_grid))
        if bitfield is not None:
            model.density_bitfield.copy_(torch.from_numpy(bitfield))
    return model.to(device).eval()
