"""Autograd bindings of the differentiable small-batch path (include/ngp_b200.h: ngp_render_diff_forward / _backward,
ngp_density_backward).

Upstream, gradients w.r.t. rays exist only on the sampler path `run()` (validate.py:117 switches cuda_ray off for the
estimator): gridencoder/grid.py:62-90, shencoder/sphere_harmonics.py:41-58 and ffmlp/ffmlp.py:51-96 return input
gradients from their backward kernels, while raymarching.py:161-230 (march_rays_train) returns sample positions with no
backward.  Like those Functions (their `@once_differentiable` is commented out), the backward passes here return plain
kernel results: a second differentiation (torch.autograd.functional.hessian in nav/estimator_helpers.py:297) sees them
as constants, exactly as it does upstream.

The weights and the occupancy grid must not change between forward and backward (the backward pass recomputes the
forward instead of storing it); the renderer registers them again before the backward launch, so a change is picked up
rather than ignored, but the gradient is then that of the new weights.
"""
import torch

from .. import _cabi


class RenderRays(torch.autograd.Function):
    """image, depth, weights_sum = RenderRays.apply(rays_o [N,3], rays_d [N,3], model, dt_gamma, max_steps, bg_color)."""

    @staticmethod
    def forward(ctx, rays_o, rays_d, model, dt_gamma, max_steps, bg_color):
        lib = _cabi.load()
        rays_o = rays_o.detach().float().contiguous()
        rays_d = rays_d.detach().float().contiguous()
        N, device = rays_o.shape[0], rays_o.device
        image = torch.empty(N, 3, dtype=torch.float32, device=device)
        depth = torch.empty(N, dtype=torch.float32, device=device)
        weights_sum = torch.empty(N, dtype=torch.float32, device=device)
        finals = torch.empty(N, 4, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            handle = model._native.get(device)
            model._set_scene(model._native, handle)
            _cabi.check(lib.ngp_render_diff_forward(handle, _cabi.ptr(rays_o), _cabi.ptr(rays_d), N, dt_gamma, max_steps, bg_color,
                                                    _cabi.ptr(image), _cabi.ptr(depth), _cabi.ptr(weights_sum), _cabi.ptr(finals), None,
                                                    _cabi.stream()), "ngp_render_diff_forward")
        ctx.save_for_backward(rays_o, rays_d, finals)
        ctx.model, ctx.args = model, (dt_gamma, max_steps, bg_color)
        ctx.mark_non_differentiable(depth)     # raymarching.cu:606-699 propagates nothing through depth either
        return image, depth, weights_sum

    @staticmethod
    def backward(ctx, g_image, g_depth, g_ws):
        lib = _cabi.load()
        rays_o, rays_d, finals = ctx.saved_tensors
        model = ctx.model
        dt_gamma, max_steps, bg_color = ctx.args
        N, device = rays_o.shape[0], rays_o.device
        g_image = (torch.zeros(N, 3, device=device) if g_image is None else g_image).detach().float().contiguous()
        g_ws = None if g_ws is None else g_ws.detach().float().contiguous()
        g_o = torch.empty(N, 3, dtype=torch.float32, device=device)
        g_d = torch.empty(N, 3, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            if not model._configure_native(model._native):
                raise RuntimeError("RenderRays.backward: the model can no longer be served by the fused path")
            handle = model._native.get(device)
            model._set_scene(model._native, handle)
            _cabi.check(lib.ngp_render_diff_backward(handle, _cabi.ptr(rays_o), _cabi.ptr(rays_d), N, dt_gamma, max_steps, bg_color,
                                                     _cabi.ptr(finals), _cabi.ptr(g_image), _cabi.ptr(g_ws), _cabi.ptr(g_o), _cabi.ptr(g_d),
                                                     _cabi.stream()), "ngp_render_diff_backward")
        return g_o, g_d, None, None, None, None


class DensityPoints(torch.autograd.Function):
    """sigma [M], geo_feat [M,15] (half, no gradient) = DensityPoints.apply(x [M,3], model)."""

    @staticmethod
    def forward(ctx, x, model):
        lib = _cabi.load()
        x = x.detach().float().contiguous()
        M, device = x.shape[0], x.device
        sigma = torch.empty(M, dtype=torch.float32, device=device)
        geo = torch.empty(M, model.geo_feat_dim, dtype=torch.float16, device=device)
        with torch.cuda.device(device):
            _cabi.check(lib.ngp_density_forward(model._native.get(device), _cabi.ptr(x), M, 0, _cabi.ptr(sigma), _cabi.ptr(geo),
                                                _cabi.stream()), "ngp_density_forward")
        ctx.save_for_backward(x)
        ctx.model = model
        ctx.mark_non_differentiable(geo)
        return sigma, geo

    @staticmethod
    def backward(ctx, g_sigma, g_geo):
        lib = _cabi.load()
        (x,) = ctx.saved_tensors
        model = ctx.model
        M, device = x.shape[0], x.device
        g_sigma = g_sigma.detach().float().contiguous()
        g_x = torch.empty(M, 3, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            if not model._configure_native(model._native):
                raise RuntimeError("DensityPoints.backward: the model can no longer be served by the fused path")
            _cabi.check(lib.ngp_density_backward(model._native.get(device), _cabi.ptr(x), _cabi.ptr(g_sigma), M, 0, _cabi.ptr(g_x),
                                                 _cabi.stream()), "ngp_density_backward")
        return g_x, None


class GetRays(torch.autograd.Function):
    """rays_o, rays_d [B,H*W,3] = GetRays.apply(poses [B,4,4], fx, fy, cx, cy, H, W): nerf/utils.py:52-116 (N <= 0) with the gradient
    w.r.t. the poses that the reference's torch sequence carries -- what Estimator.measurement_fn differentiates through
    (nav/estimator_helpers.py:203-210).  Forward: one kernel (ngp_get_rays).  Backward (two tiny reductions, left to torch):
    rays_o = t (broadcast) -> d t = sum over pixels of g_o;  rays_d = R dir_cam -> d R = sum over pixels of g_d (x) dir_cam."""

    @staticmethod
    def forward(ctx, poses, fx, fy, cx, cy, H, W):
        lib = _cabi.load()
        poses = poses.detach().to(torch.float32).contiguous()
        B = poses.shape[0]
        rays_o = torch.empty(B, H * W, 3, dtype=torch.float32, device=poses.device)
        rays_d = torch.empty(B, H * W, 3, dtype=torch.float32, device=poses.device)
        with torch.cuda.device(poses.device):
            _cabi.check(lib.ngp_get_rays(_cabi.ptr(poses), B, H, W, fx, fy, cx, cy, _cabi.ptr(rays_o), _cabi.ptr(rays_d), _cabi.stream()),
                        "ngp_get_rays")
        ctx.args = (fx, fy, cx, cy, H, W, poses.device)
        return rays_o, rays_d

    @staticmethod
    def backward(ctx, g_o, g_d):
        fx, fy, cx, cy, H, W, device = ctx.args
        B = (g_o if g_o is not None else g_d).shape[0]
        g_pose = torch.zeros(B, 4, 4, dtype=torch.float32, device=device)
        if g_o is not None:
            g_pose[:, :3, 3] = g_o.sum(1)
        if g_d is not None:
            j, i = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float32) + 0.5,
                                  torch.arange(W, device=device, dtype=torch.float32) + 0.5, indexing="ij")
            dirs = torch.stack(((i - cx) / fx, (j - cy) / fy, torch.ones_like(i)), -1).reshape(H * W, 3)
            dirs = dirs / dirs.norm(dim=-1, keepdim=True)
            g_pose[:, :3, :3] = torch.einsum("bpk,pc->bkc", g_d.float(), dirs)      # rays_d[b,p,k] = sum_c R[b,k,c] dirs[p,c]
        return g_pose, None, None, None, None, None, None
