"""fp64 differentiable restatement of the CUDA-ray render for fixed sample parameters t (TEST INFRASTRUCTURE ONLY).

The reference has no differentiable CUDA-ray render (raymarching.py:161-230: march_rays_train has no backward), so the derivative the
B200 kernels must produce is DEFINED here, from the reference's own formulas, in torch float64 with autograd:

    x = clamp(o + t d)                          raymarching.cu:756-758, t constant
    enc = trilinear hash grid (fp16 table)      gridencoder.cu:75-176 (indices :35-72), u = (x + bound) / (2 bound) grid.py:144
    h = sigma_net(enc); sigma = exp(h0) * density_scale; geo = h[1:16]     network_ff.py:56-61, renderer.py:365
    rgb = sigmoid(color_net([SH4(d), geo, 0]))  network_ff.py:64-72, shencoder.cu:44-69
    composite front to back, stop after the first sample whose incoming transmittance is <= 1e-4      raymarching.cu:866-894
    image = sum w c + (1 - sum w) bg            renderer.py:374

No fp16 rounding of activations (the table and weights ARE the fp16 values): the forward agrees with the C oracle / the kernels to
~1e-3, and its autograd derivative is what tests/test_gpu_diff.py compares the kernels' gradients with.  tests/test_diff_oracle.py pins
this file on the CPU: forward against oracle/ngp_oracle.c, derivative against central differences of itself.
"""
import numpy as np
import torch

P1, P2 = 2654435761, 805459861


def level_table(weights):
    """Per level (scale, stride, size, offset, hashed) exactly as ngp_ctx_set_network derives them (gridencoder.cu:60-69,126-127)."""
    offsets = np.asarray(weights["offsets"], np.int64)
    S, H = np.float32(weights["S"]), np.float32(weights["H"])
    out = []
    for l in range(len(offsets) - 1):
        scale = np.float32(np.exp2(np.float32(l) * S)) * H - np.float32(1.0)
        res = int(np.ceil(scale)) + 1
        stride, size = res + 1, int(offsets[l + 1] - offsets[l])
        s = 1
        for _ in range(3):
            if s <= size:
                s *= stride
        out.append(dict(scale=float(scale), stride=stride, size=size, offset=int(offsets[l]), hashed=s > size))
    return out


def grid_encode(u, table, levels):
    """u [M,3] float64 in [0,1] (rows outside -> zeros, gridencoder.cu:99-123), table [rows,2] float64 -> [M, 2 L]."""
    inside = ((u >= 0) & (u <= 1)).all(-1, keepdim=True)
    feats = []
    for lv in levels:
        p = u * lv["scale"] + 0.5
        g = torch.floor(p.detach())
        f = p - g
        g = g.long().clamp_(min=0)
        acc = 0
        for k in range(8):
            bit = [(k >> d) & 1 for d in range(3)]
            c = [g[:, d] + bit[d] for d in range(3)]
            if lv["hashed"]:
                idx = ((c[0] & 0xffffffff) ^ ((c[1] * P1) & 0xffffffff) ^ ((c[2] * P2) & 0xffffffff)) % lv["size"]
            else:
                idx = (c[0] + c[1] * lv["stride"] + c[2] * lv["stride"] * lv["stride"]) % lv["size"]
            w = 1
            for d in range(3):
                w = w * (f[:, d] if bit[d] else 1 - f[:, d])
            acc = acc + w[:, None] * table[lv["offset"] + idx]
        feats.append(acc)
    return torch.where(inside, torch.cat(feats, -1), torch.zeros((), dtype=u.dtype, device=u.device))


def sh4(d):
    x, y, z = d.unbind(-1)
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    return torch.stack([
        torch.full_like(x, 0.28209479177387814), -0.48860251190291987 * y, 0.48860251190291987 * z, -0.48860251190291987 * x,
        1.0925484305920792 * xy, -1.0925484305920792 * yz, 0.94617469575755997 * z2 - 0.31539156525251999, -1.0925484305920792 * xz,
        0.54627421529603959 * x2 - 0.54627421529603959 * y2, 0.59004358992664352 * y * (-3.0 * x2 + y2), 2.8906114426405538 * xy * z,
        0.45704579946446572 * y * (1.0 - 5.0 * z2), 0.3731763325901154 * z * (5.0 * z2 - 3.0), 0.45704579946446572 * x * (1.0 - 5.0 * z2),
        1.4453057213202769 * z * (x2 - y2), 0.59004358992664352 * x * (-x2 + 3.0 * y2)], -1)


def mlp(x, w_flat, in_dim, num_layers):
    """ffmlp.py:121 layout: [64][in] + (num_layers - 1) x [64][64] + [16][64]; ReLU between, none at the end."""
    w, off = w_flat, 0
    W = w[off:off + 64 * in_dim].view(64, in_dim); off += 64 * in_dim
    x = torch.relu(x @ W.T)
    for _ in range(num_layers - 1):
        W = w[off:off + 4096].view(64, 64); off += 4096
        x = torch.relu(x @ W.T)
    W = w[off:off + 1024].view(16, 64)
    return x @ W.T


class Net:
    def __init__(self, weights, device="cpu"):
        f64 = dict(dtype=torch.float64, device=device)
        self.table = torch.from_numpy(np.asarray(weights["table"], np.float16).astype(np.float64)).to(**f64)
        self.w_sigma = torch.from_numpy(np.asarray(weights["w_sigma"], np.float16).astype(np.float64)).to(**f64)
        self.w_color = torch.from_numpy(np.asarray(weights["w_color"], np.float16).astype(np.float64)).to(**f64)
        self.levels = level_table(weights)
        self.bound, self.density_scale = float(weights["bound"]), float(weights["density_scale"])
        self.nl, self.nlc = int(weights["num_layers"]), int(weights["num_layers_color"])

    def density(self, x):
        """sigma (without density_scale, like NeRFNetwork.density) and h."""
        h = mlp(grid_encode((x + self.bound) / (2 * self.bound), self.table, self.levels), self.w_sigma, 32, self.nl)
        return torch.exp(h[:, 0]), h

    def forward(self, x, d):
        sigma, h = self.density(x)
        hin = torch.cat([sh4(d), h[:, 1:], torch.zeros_like(h[:, :1])], -1)
        rgb = torch.sigmoid(mlp(hin, self.w_color, 32, self.nlc)[:, :3])
        return sigma, rgb


def render(net, rays_o, rays_d, t, dt, count, bg=1.0):
    """rays_o, rays_d [N,3] float64 (requires_grad allowed); t, dt [N,S] sample parameters / step sizes (constants), count [N] valid
    samples per ray.  Returns image [N,3], weights_sum [N]."""
    N, S = t.shape
    valid = torch.arange(S, device=t.device)[None, :] < count[:, None]
    x = (rays_o[:, None, :] + t[..., None] * rays_d[:, None, :]).clamp(-net.bound, net.bound)
    d = rays_d[:, None, :].expand(N, S, 3)
    sigma, rgb = net.forward(x.reshape(-1, 3), d.reshape(-1, 3))
    sigma = (sigma * net.density_scale).view(N, S) * valid
    rgb = rgb.view(N, S, 3)
    alpha = 1 - torch.exp(-sigma * dt)
    T = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha[:, :-1]], 1), 1)        # incoming transmittance
    stopped = (torch.cumsum((T.detach() <= 1e-4).long(), 1) - (T.detach() <= 1e-4).long()) > 0     # an EARLIER sample saw T <= 1e-4
    w = alpha * T * (valid & ~stopped)
    ws = w.sum(1)
    image = (w[..., None] * rgb).sum(1) + (1 - ws)[:, None] * bg
    return image, ws


def samples_from_deltas(deltas, nears, N, S):
    """deltas [N*S(+pad), 2] as march_rays writes them for ONE chunk of n_step = S samples per ray starting at t = near
    (raymarching.cu:754-812: column 0 = dt, column 1 = t_new - t_previous; rows past a ray's last sample stay zero).
    Returns float64 (t [N,S], dt [N,S], count [N]); t = the parameter the sample was taken at."""
    dl = torch.as_tensor(deltas)[: N * S].double().view(N, S, 2)
    count = (dl[..., 0] != 0).sum(1)
    t_after = torch.as_tensor(nears).double()[:, None] + torch.cumsum(dl[..., 1], 1)
    return t_after - dl[..., 0], dl[..., 0], count
