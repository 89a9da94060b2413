"""CPU restatement (torch, fp32) of the reference's pure-PyTorch renderer `NeRFRenderer.run` -- the `cuda_ray=False` path.

TEST INFRASTRUCTURE / CPU BASELINE ONLY -- imported by tests/ and by bench.py's cpu legs, never by nerfsafetyvalidation_b200/
(whose `NeRFRenderer.run` raises: the product has no CPU or PyTorch path).

This is BASELINE.json configs[0]: the vanilla NeRFNetwork (frequency encoders, nn.Linear MLPs) rendering with the hierarchical
64 + 64 sampler on the host cores.  What follows what:
    freq_encode      encoding.py:5-43          FreqEncoder(max_freq_log2 = 5, N_freqs = 6): [x, sin(x f0), cos(x f0), ...]
                                               (== freqencoder/src/freqencoder.cu:48-57 ordering, the module `get_encoder`
                                               instantiates at encoding.py:57)
    VanillaNeRF      nerf/network.py:10-191    sigma_net 39 -> 64 -> 16, color_net 54 -> 64 -> 64 -> 3, no biases; 11 264 params
    sample_pdf       nerf/renderer.py:12-46    inverse-CDF resampling, det = True in eval mode
    run              nerf/renderer.py:125-258  uniform samples in [near, far] -> density -> weights -> sample_pdf -> density of the
                                               new points -> sort / gather -> weights -> masked colour -> composite
    render           nerf/renderer.py:549-588  staged: chunks of max_ray_batch rays
near / far come from the C oracle (oracle/ngp_oracle.c:near_far, pinned bitwise against the reference's CUDA kernel).
Pinned against the imported reference itself: tests/golden/pyref_cfg1.npz (tests/golden/make_golden_pyref.py).
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import oracle as orc


def freq_encode(x, n_freqs=6):
    out = [x]
    for k in range(n_freqs):
        f = float(2.0 ** k)   # 2 ** linspace(0, n_freqs - 1, n_freqs)
        out.append(torch.sin(x * f))
        out.append(torch.cos(x * f))
    return torch.cat(out, dim=-1)


class VanillaNeRF:
    """Weights only; creation order and initialiser are nn.Linear's, so torch.manual_seed(s) gives the reference's weights."""

    def __init__(self, seed=0, num_layers=2, hidden=64, geo=15, num_layers_color=3, bound=1.0, min_near=0.2, density_scale=1.0):
        torch.manual_seed(seed)
        enc = 3 + 3 * 6 * 2
        dims_s = [enc] + [hidden] * (num_layers - 1) + [1 + geo]
        dims_c = [enc + geo] + [hidden] * (num_layers_color - 1) + [3]
        self.sigma = [torch.nn.Linear(a, b, bias=False).weight.detach() for a, b in zip(dims_s[:-1], dims_s[1:])]
        self.color = [torch.nn.Linear(a, b, bias=False).weight.detach() for a, b in zip(dims_c[:-1], dims_c[1:])]
        self.bound, self.min_near, self.density_scale = float(bound), float(min_near), float(density_scale)

    def n_params(self):
        return sum(w.numel() for w in self.sigma + self.color)

    def load(self, sigma_w, color_w):
        self.sigma = [torch.as_tensor(w, dtype=torch.float32) for w in sigma_w]
        self.color = [torch.as_tensor(w, dtype=torch.float32) for w in color_w]
        return self

    def density(self, x):   # network.py:126-143
        h = freq_encode(x)
        for k, w in enumerate(self.sigma):
            h = F.linear(h, w)
            if k != len(self.sigma) - 1:
                h = F.relu(h)
        return torch.exp(h[..., 0]), h[..., 1:]

    def rgb(self, d, geo, mask):   # network.py:162-191 (masked inference)
        out = torch.zeros(mask.shape[0], 3, dtype=d.dtype)
        if not bool(mask.any()):
            return out
        h = torch.cat([freq_encode(d[mask]), geo[mask]], dim=-1)
        for k, w in enumerate(self.color):
            h = F.linear(h, w)
            if k != len(self.color) - 1:
                h = F.relu(h)
        out[mask] = torch.sigmoid(h)
        return out


def sample_pdf(bins, weights, n_samples):
    """renderer.py:12-46 with det=True.  bins [B,T], weights [B,T-1] -> [B,n_samples]."""
    weights = weights + 1e-5
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    u = torch.linspace(0.0 + 0.5 / n_samples, 1.0 - 0.5 / n_samples, steps=n_samples).expand(list(cdf.shape[:-1]) + [n_samples]).contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.clamp(inds - 1, min=0)
    above = torch.clamp(inds, max=cdf.shape[-1] - 1)
    cdf_lo, cdf_hi = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    bin_lo, bin_hi = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = cdf_hi - cdf_lo
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    return bin_lo + (u - cdf_lo) / denom * (bin_hi - bin_lo)


def _weights(z, sigma, sample_dist, density_scale):   # renderer.py:184-189 == :222-226
    deltas = torch.cat([z[..., 1:] - z[..., :-1], sample_dist * torch.ones_like(z[..., :1])], dim=-1)
    alphas = 1 - torch.exp(-deltas * density_scale * sigma)
    shifted = torch.cat([torch.ones_like(alphas[..., :1]), 1 - alphas + 1e-15], dim=-1)
    return alphas * torch.cumprod(shifted, dim=-1)[..., :-1]


@torch.no_grad()
def run(net, rays_o, rays_d, num_steps=64, upsample_steps=64, bg_color=1.0):
    """renderer.py:125-258, eval mode, perturb=False.  rays [N,3] fp32 tensors -> dict of image [N,3], depth, weights_sum,
    aggregated_density [N]."""
    N = rays_o.shape[0]
    b = net.bound
    aabb = np.array([-b, -b, -b, b, b, b], np.float32)
    nears, fars = orc.near_far_from_aabb(rays_o.numpy(), rays_d.numpy(), aabb, net.min_near)
    nears, fars = torch.from_numpy(nears)[:, None], torch.from_numpy(fars)[:, None]
    lo, hi = torch.from_numpy(aabb[:3]), torch.from_numpy(aabb[3:])
    z = nears + (fars - nears) * torch.linspace(0.0, 1.0, num_steps)[None].expand(N, num_steps)
    sample_dist = (fars - nears) / num_steps
    xyz = torch.min(torch.max(rays_o[:, None] + rays_d[:, None] * z[..., None], lo), hi)
    sigma, geo = net.density(xyz.reshape(-1, 3))
    sigma, geo = sigma.view(N, num_steps), geo.view(N, num_steps, -1)
    if upsample_steps > 0:
        deltas = torch.cat([z[..., 1:] - z[..., :-1], sample_dist * torch.ones_like(z[..., :1])], dim=-1)
        w = _weights(z, sigma, sample_dist, net.density_scale)
        mid = z[..., :-1] + 0.5 * deltas[..., :-1]
        new_z = sample_pdf(mid, w[:, 1:-1], upsample_steps)
        new_xyz = torch.min(torch.max(rays_o[:, None] + rays_d[:, None] * new_z[..., None], lo), hi)
        new_sigma, new_geo = net.density(new_xyz.reshape(-1, 3))
        z, order = torch.sort(torch.cat([z, new_z], dim=1), dim=1)
        xyz = torch.gather(torch.cat([xyz, new_xyz], dim=1), 1, order[..., None].expand(-1, -1, 3))
        sigma = torch.gather(torch.cat([sigma, new_sigma.view(N, upsample_steps)], dim=1), 1, order)
        geo_all = torch.cat([geo, new_geo.view(N, upsample_steps, -1)], dim=1)
        geo = torch.gather(geo_all, 1, order[..., None].expand_as(geo_all))
    w = _weights(z, sigma, sample_dist, net.density_scale)
    T = z.shape[1]
    dirs = rays_d[:, None].expand(N, T, 3)
    rgbs = net.rgb(dirs.reshape(-1, 3), geo.reshape(N * T, -1), (w > 1e-4).reshape(-1)).view(N, T, 3)
    weights_sum = w.sum(dim=-1)
    depth = torch.sum(w * ((z - nears) / (fars - nears)).clamp(0, 1), dim=-1)
    image = torch.sum(w[..., None] * rgbs, dim=-2) + (1 - weights_sum)[..., None] * bg_color
    return {"image": image, "depth": depth, "weights_sum": weights_sum, "aggregated_density": torch.sum(w * sigma, dim=1)}


@torch.no_grad()
def render(net, rays_o, rays_d, max_ray_batch=4096, **kw):
    """renderer.py:549-588 with staged=True: chunks of max_ray_batch rays through run()."""
    N = rays_o.shape[0]
    outs = [run(net, rays_o[h:h + max_ray_batch], rays_d[h:h + max_ray_batch], **kw) for h in range(0, N, max_ray_batch)]
    return {k: torch.cat([o[k] for o in outs]) for k in outs[0]}


def get_rays(poses, fx, fy, cx, cy, H, W):
    """nerf/utils.py:52-116 (N = -1) in numpy fp32 with the operation order of csrc/ops_raymarch.cu:k_get_rays
    (IEEE mul / add / div / sqrt, no contraction): poses [F,4,4] -> rays_o, rays_d [F, H*W, 3]."""
    f32 = np.float32
    poses = np.asarray(poses, f32)
    p = np.arange(H * W)
    i = (p % W).astype(f32) + f32(0.5)
    j = (p // W).astype(f32) + f32(0.5)
    x = (i - f32(cx)) / f32(fx)
    y = (j - f32(cy)) / f32(fy)
    nrm = np.sqrt((x * x + y * y) + f32(1.0))
    d = np.stack([x / nrm, y / nrm, f32(1.0) / nrm], -1)                 # [HW,3]
    R = poses[:, None, :3, :3]                                           # [F,1,3(k),3]
    rays_d = (d[None, :, None, 0] * R[..., 0] + d[None, :, None, 1] * R[..., 1]) + d[None, :, None, 2] * R[..., 2]
    rays_o = np.broadcast_to(poses[:, None, :3, 3], rays_d.shape)
    return np.ascontiguousarray(rays_o, f32), np.ascontiguousarray(rays_d, f32)
