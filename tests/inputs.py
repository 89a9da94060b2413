"""Seeded input generators shared by the GPU parity tests, the golden-vector script and the CPU oracle tests.
torch CPU generators only, so every machine sees the same bytes."""
import torch


def make_rays(n, seed, device="cpu"):
    """inside/outside origins, axis-parallel directions (d = +-0 -> rcp = +-inf), grazing and missing rays."""
    g = torch.Generator().manual_seed(seed)
    o = (torch.rand(n, 3, generator=g) * 4 - 2)
    o[: n // 4] = torch.rand(n // 4, 3, generator=g) * 1.8 - 0.9           # inside the box
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    k = n // 16
    d[:k, 0] = 0.0
    d[k:2 * k, 1] = -0.0
    d[2 * k:3 * k] = torch.tensor([0.0, 0.0, 1.0])
    d[3 * k:4 * k] = torch.tensor([-0.0, 1.0, -0.0])
    o[4 * k:5 * k, 0] = 1.0                                                # origin on a face
    d[5 * k:6 * k] = -o[5 * k:6 * k] / o[5 * k:6 * k].norm(dim=-1, keepdim=True)   # aimed at the centre
    return o.to(device), d.contiguous().to(device)


def grid_inputs(n, seed, device="cpu"):
    """uniform u in [0,1], exact 0 and 1, out-of-bounds rows, values on cell boundaries k/res."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, 3, generator=g)
    x[:64] = 0.0
    x[64:128] = 1.0
    x[128:192, 0] = 1.5          # out of bounds -> zeros
    x[192:256, 2] = -0.25
    res = torch.tensor([16, 23, 31, 43, 59, 81, 112, 154, 213, 295, 407, 562, 777, 1073, 1483, 2048], dtype=torch.float32)
    k = torch.randint(0, 16, (1024,), generator=g)
    x[256:1280, 0] = torch.floor(torch.rand(1024, generator=g) * res[k]) / res[k]   # on cell boundaries
    return x.to(device)


def unit_dirs(n, seed, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    d = torch.randn(n, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    return d.contiguous().to(device)
