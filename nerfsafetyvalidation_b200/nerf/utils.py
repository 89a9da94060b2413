"""The two helpers of the reference's nerf/utils.py that sit on the render path: `custom_meshgrid`, `get_rays`."""
import torch

from .. import _cabi


def custom_meshgrid(*args):
    """nerf/utils.py:33-38"""
    return torch.meshgrid(*args, indexing="ij")


def get_rays(poses, intrinsics, H, W, N=-1, error_map=None):
    """nerf/utils.py:52-116 for the inference case (N <= 0: every pixel, pixel centres at +0.5).
    poses [B,4,4] cam2world (CUDA), intrinsics (fx, fy, cx, cy) -> dict(rays_o [B,H*W,3], rays_d [B,H*W,3]).
    One kernel (ngp_get_rays) instead of the reference's meshgrid / stack / norm / matmul sequence; no CPU fallback.  Poses that
    require grad keep their graph (nerf/autograd.py:GetRays), as they do through the reference's torch sequence."""
    if N > 0:
        raise NotImplementedError("get_rays: random pixel sub-sampling (N > 0) is the training path")
    _cabi.require_cuda(poses)
    fx, fy, cx, cy = (float(v) for v in intrinsics)
    if torch.is_grad_enabled() and poses.requires_grad:
        # the estimator differentiates its loss w.r.t. the pose through this call (nav/estimator_helpers.py:203-210)
        from .autograd import GetRays
        if tuple(poses.shape[1:]) != (4, 4):
            raise RuntimeError(f"get_rays: poses must be [B,4,4], got {tuple(poses.shape)}")
        rays_o, rays_d = GetRays.apply(poses, fx, fy, cx, cy, int(H), int(W))
        return {"rays_o": rays_o, "rays_d": rays_d}
    poses = poses.detach().to(torch.float32).contiguous()
    B = poses.shape[0]
    if tuple(poses.shape[1:]) != (4, 4):
        raise RuntimeError(f"get_rays: poses must be [B,4,4], got {tuple(poses.shape)}")
    rays_o = torch.empty(B, H * W, 3, dtype=torch.float32, device=poses.device)
    rays_d = torch.empty(B, H * W, 3, dtype=torch.float32, device=poses.device)
    with torch.cuda.device(poses.device):
        _cabi.check(_cabi.load().ngp_get_rays(_cabi.ptr(poses), B, H, W, fx, fy, cx, cy, _cabi.ptr(rays_o), _cabi.ptr(rays_d), _cabi.stream()),
                    "ngp_get_rays")
    return {"rays_o": rays_o, "rays_d": rays_d}
