"""The two helpers of the reference's nerf/utils.py that sit on the render path: `custom_meshgrid`, `get_rays`."""
import torch


def custom_meshgrid(*args):
    """nerf/utils.py:33-38"""
    return torch.meshgrid(*args, indexing="ij")


@torch.no_grad()
def get_rays(poses, intrinsics, H, W, N=-1, error_map=None):
    """nerf/utils.py:52-116 for the inference case (N <= 0: every pixel, pixel centres at +0.5).
    poses [B,4,4] cam2world, intrinsics (fx, fy, cx, cy) -> dict(rays_o [B,H*W,3], rays_d [B,H*W,3])."""
    if N > 0:
        raise NotImplementedError("get_rays: random pixel sub-sampling (N > 0) is the training path")
    device = poses.device
    B = poses.shape[0]
    fx, fy, cx, cy = intrinsics
    i, j = custom_meshgrid(torch.linspace(0, W - 1, W, device=device), torch.linspace(0, H - 1, H, device=device))
    i = i.t().reshape([1, H * W]).expand([B, H * W]) + 0.5
    j = j.t().reshape([1, H * W]).expand([B, H * W]) + 0.5
    zs = torch.ones_like(i)
    xs = (i - cx) / fx * zs
    ys = (j - cy) / fy * zs
    directions = torch.stack((xs, ys, zs), dim=-1)
    directions = directions / torch.norm(directions, dim=-1, keepdim=True)
    rays_d = directions @ poses[:, :3, :3].transpose(-1, -2)
    rays_o = poses[..., :3, 3]
    rays_o = rays_o[..., None, :].expand_as(rays_d)
    return {"rays_o": rays_o, "rays_d": rays_d}
