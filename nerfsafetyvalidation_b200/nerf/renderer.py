"""Drop-in for the inference side of the reference's nerf/renderer.py: `NeRFRenderer`.

`render(rays_o, rays_d, staged=False, max_ray_batch=4096, **kwargs)` and `run_cuda(...)` keep the reference's
signatures and result dict (nerf/renderer.py:261-386, 549-588).  Two execution modes for the inference branch:

  fused (default)   the whole loop of renderer.py:329-376 is ONE persistent cooperative kernel (ngp_render):
                    near/far -> {march -> hash-grid -> MLPs -> composite -> compaction (survivors appended unordered;
                    per-ray results do not depend on the order)} with the reference's
                    global `n_step = clamp(N // n_alive, 1, 8)` schedule evaluated on the device, then background
                    mix and depth normalisation.  No host sync, no intermediate tensors in HBM.
  stepwise          the reference's Python loop verbatim, on our stand-alone ops (one launch per op per
                    iteration); kept for kernel-boundary parity tests (`fused=False`).

Training raises.  `run()` (cuda_ray=False, what `validate.py -O` selects) is the reference's hierarchical sampler on the device,
with our kernels for near/far and every network query; it is a convenience for callers of that flag, not the benchmarked path.
"""
import ctypes as C
import math

import numpy as np
import torch
import torch.nn as nn

from .. import _cabi
from .. import raymarching


class _NativeCtx:
    """Owner of one ngp_ctx.  Not copied by deepcopy (NerfSimulator.py:37 deep-copies the model): the copy
    lazily creates its own context on first use."""

    def __init__(self):
        self.handle = None
        self.device = None    # CUDA device index the context (and its workspace) lives on
        self.keys = None      # what the context was last configured with
        self.keep = None      # tensors the context borrows
        self.options = {}     # ngp_ctx_set_option values, re-applied when the context is re-created

    def get(self, device=None):
        """The context for `device` (a torch.device / index; default: the current device).  A context is bound to the device it
        was created on (its workspace is allocated there), so a model that moved gets a new one."""
        import torch
        dev = torch.cuda.current_device() if device is None else torch.device(device).index
        if dev is None:
            dev = torch.cuda.current_device()
        if self.handle is not None and self.device != dev:
            self.close()
        if self.handle is None:
            lib = _cabi.load()
            h = C.c_void_p()
            with torch.cuda.device(dev):
                _cabi.check(lib.ngp_ctx_create(C.byref(h)), "ngp_ctx_create")
            self.handle, self.device, self.keys, self.keep = h, dev, None, None
            for k, v in self.options.items():
                _cabi.check(lib.ngp_ctx_set_option(h, k.encode(), int(v)), "ngp_ctx_set_option")
        return self.handle

    def set_option(self, name, value):
        """Epoch-policy tunables of the fused renderer (include/ngp_b200.h: ngp_ctx_set_option); no option changes any output."""
        self.options[name] = int(value)
        if self.handle is not None:
            _cabi.check(_cabi.load().ngp_ctx_set_option(self.handle, name.encode(), int(value)), "ngp_ctx_set_option")

    def close(self):
        if self.handle is not None:
            _cabi.load().ngp_ctx_destroy(self.handle)
        self.handle, self.device, self.keys, self.keep = None, None, None, None

    def __deepcopy__(self, memo):
        return _NativeCtx()

    def __getstate__(self):
        return {}

    def __setstate__(self, state):
        self.handle, self.device, self.keys, self.keep, self.options = None, None, None, None, {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NeRFRenderer(nn.Module):
    def __init__(self, bound=1, cuda_ray=False, density_scale=1, min_near=0.2, density_thresh=0.01, bg_radius=-1):
        super().__init__()
        self.bound = bound
        self.cascade = 1 + math.ceil(math.log2(bound))
        self.grid_size = 128
        self.density_scale = density_scale
        self.min_near = min_near
        self.density_thresh = density_thresh
        self.bg_radius = bg_radius

        aabb_train = torch.tensor([-bound, -bound, -bound, bound, bound, bound], dtype=torch.float32)
        self.register_buffer("aabb_train", aabb_train)
        self.register_buffer("aabb_infer", aabb_train.clone())

        self.cuda_ray = cuda_ray
        if cuda_ray:
            self.register_buffer("density_grid", torch.zeros([self.cascade, self.grid_size ** 3]))
            self.register_buffer("density_bitfield", torch.zeros(self.cascade * self.grid_size ** 3 // 8, dtype=torch.uint8))
            self.mean_density = 0
            self.iter_density = 0
            self.register_buffer("step_counter", torch.zeros(16, 2, dtype=torch.int32))
            self.mean_count = 0
            self.local_step = 0
        self._native = _NativeCtx()
        self.fused = True   # run_cuda default mode
        # Opt-in: inference batches of at most this many rays go to the warp-per-ray kernel (ngp_render_diff_forward) instead of the
        # persistent one, whose fixed cost (a cooperative launch, two grid barriers per epoch) dominates small batches: 1024 rays
        # 0.13 ms instead of 1.5 ms, break-even around 40 000 rays (profiles/r02i_small_batch.json).  Off by default because the two
        # differ where the reference itself depends on the batch: its marcher restarts every chunk of n_step = clamp(N // n_alive, 1, 8)
        # samples from a re-summed t (raymarching.cu:723-725, 870, 905), which after a large empty gap can be one ulp off and move one
        # sample across a cell face.  The persistent kernel reproduces that schedule exactly; the warp-per-ray kernel marches each ray
        # as ONE chunk (measured: 2 of 65 536 rays take one sample more / fewer, image difference 1.4e-3 on those, <= 1e-6 elsewhere).
        self.small_batch_rays = 0

    def forward(self, x, d):
        raise NotImplementedError()

    def density(self, x):
        raise NotImplementedError()

    def color(self, x, d, mask=None, **kwargs):
        raise NotImplementedError()

    def reset_extra_state(self):
        if not self.cuda_ray:
            return
        self.density_grid.zero_()
        self.mean_density = 0
        self.iter_density = 0
        self.step_counter.zero_()
        self.mean_count = 0
        self.local_step = 0

    @staticmethod
    def _sample_pdf(bins, weights, n_samples, det):
        """nerf/renderer.py:12-46: inverse-CDF resampling.  bins [B,T], weights [B,T-1] -> [B,n_samples]."""
        weights = weights + 1e-5
        cdf = torch.cumsum(weights / torch.sum(weights, -1, keepdim=True), -1)
        cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
        shape = list(cdf.shape[:-1]) + [n_samples]
        if det:
            u = torch.linspace(0.5 / n_samples, 1.0 - 0.5 / n_samples, steps=n_samples, device=cdf.device, dtype=cdf.dtype).expand(shape)
        else:
            u = torch.rand(shape, device=cdf.device, dtype=cdf.dtype)
        u = u.contiguous()
        inds = torch.searchsorted(cdf, u, right=True)
        lo, hi = torch.clamp(inds - 1, min=0), torch.clamp(inds, max=cdf.shape[-1] - 1)
        cdf_lo, cdf_hi = torch.gather(cdf, 1, lo), torch.gather(cdf, 1, hi)
        bin_lo, bin_hi = torch.gather(bins, 1, lo), torch.gather(bins, 1, hi)
        denom = cdf_hi - cdf_lo
        denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
        return bin_lo + (u - cdf_lo) / denom * (bin_hi - bin_lo)

    def _sample_weights(self, z, sigma, sample_dist):
        """renderer.py:184-189 == :222-226: alpha from sigma * delta, transmittance by cumprod."""
        deltas = torch.cat([z[..., 1:] - z[..., :-1], sample_dist * torch.ones_like(z[..., :1])], dim=-1)
        alphas = 1 - torch.exp(-deltas * self.density_scale * sigma)
        shifted = torch.cat([torch.ones_like(alphas[..., :1]), 1 - alphas + 1e-15], dim=-1)
        return alphas * torch.cumprod(shifted, dim=-1)[..., :-1]

    def run(self, rays_o, rays_d, num_steps=128, upsample_steps=128, bg_color=None, perturb=False, **kwargs):
        """nerf/renderer.py:125-258 -- the `cuda_ray=False` hierarchical sampler (what `validate.py -O` selects), on the DEVICE:
        near/far, every density query (ngp_density_forward) and every colour query (SH + FFMLP ops) are our kernels; torch
        only does the sampler's bookkeeping (cumprod / searchsorted / sort / gather) on device tensors.  Same result dict as the
        reference, including `aggregated_density`.  Inference only; no CPU path (CPU rays are an error)."""
        if self.training:
            raise NotImplementedError("run: only the inference branch is built (model.eval())")
        if self.bg_radius > 0:
            raise NotImplementedError("run: bg_radius > 0 (background network) is not on the validate.py path (bg_radius=-1)")
        _cabi.require_cuda(rays_o, rays_d)
        prefix = rays_o.shape[:-1]
        rays_o = rays_o.contiguous().view(-1, 3).float()
        rays_d = rays_d.contiguous().view(-1, 3).float()
        N = rays_o.shape[0]
        device = rays_o.device
        aabb = self.aabb_infer
        with torch.no_grad():
            nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, aabb, self.min_near)
            nears, fars = nears.unsqueeze(-1), fars.unsqueeze(-1)
            z = nears + (fars - nears) * torch.linspace(0.0, 1.0, num_steps, device=device).unsqueeze(0).expand(N, num_steps)
            sample_dist = (fars - nears) / num_steps
            if perturb:
                z = z + (torch.rand(z.shape, device=device) - 0.5) * sample_dist
            xyz = torch.min(torch.max(rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * z.unsqueeze(-1), aabb[:3]), aabb[3:])
            dens = {k: v.reshape(N, num_steps, -1) for k, v in self.density(xyz.reshape(-1, 3)).items()}
            if upsample_steps > 0:
                sigma = dens["sigma"].squeeze(-1).float()
                w = self._sample_weights(z, sigma, sample_dist)
                deltas = z[..., 1:] - z[..., :-1]
                new_z = self._sample_pdf(z[..., :-1] + 0.5 * deltas, w[:, 1:-1], upsample_steps, det=not self.training)
                new_xyz = torch.min(torch.max(rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * new_z.unsqueeze(-1), aabb[:3]), aabb[3:])
                new_dens = {k: v.reshape(N, upsample_steps, -1) for k, v in self.density(new_xyz.reshape(-1, 3)).items()}
                z, order = torch.sort(torch.cat([z, new_z], dim=1), dim=1)
                xyz = torch.gather(torch.cat([xyz, new_xyz], dim=1), 1, order.unsqueeze(-1).expand(-1, -1, 3))
                for k in dens:
                    both = torch.cat([dens[k], new_dens[k]], dim=1)
                    dens[k] = torch.gather(both, 1, order.unsqueeze(-1).expand_as(both))
            sigma = dens["sigma"].squeeze(-1).float()
            w = self._sample_weights(z, sigma, sample_dist)
            T = z.shape[1]
            dirs = rays_d.view(-1, 1, 3).expand(N, T, 3)
            flat = {k: v.reshape(N * T, v.shape[-1]) for k, v in dens.items()}
            mask = w > 1e-4                                                       # renderer.py:232 "hard coded"
            rgbs = self.color(xyz.reshape(-1, 3), dirs.reshape(-1, 3), mask=mask.reshape(-1), **flat).view(N, T, 3).float()
            weights_sum = w.sum(dim=-1)
            depth = torch.sum(w * ((z - nears) / (fars - nears)).clamp(0, 1), dim=-1)
            image = torch.sum(w.unsqueeze(-1) * rgbs, dim=-2)
            if bg_color is None:
                bg_color = 1
            image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
            aggregated_density = torch.sum(w * sigma, dim=1)
        return {"depth": depth.view(*prefix), "image": image.view(*prefix, 3), "weights_sum": weights_sum, "rgbs": rgbs,
                "sigmas": flat["sigma"], "aggregated_density": aggregated_density.view(*prefix)}

    # ------------------------------------------------------------------ fused-path plumbing (subclass provides the network)
    def _configure_native(self, ctx):
        """Subclasses register their network with the context; returns False if the fused path cannot serve this model."""
        return False

    def _set_scene(self, ctx_owner, handle):
        lib = _cabi.load()
        if getattr(self, "_frozen", False) and ctx_owner.keys is not None and "scene" in ctx_owner.keys:
            return   # inside `with model.frozen_weights():` the scene registered by the first call stands (no aabb read-back per call)
        aabb = self.aabb_train if self.training else self.aabb_infer
        bf = self.density_bitfield
        key = (bf.data_ptr(), self.cascade, self.grid_size, tuple(aabb.tolist()), float(self.min_near))
        if ctx_owner.keys is None or ctx_owner.keys.get("scene") != key:
            arr = (C.c_float * 6)(*aabb.tolist())
            _cabi.check(lib.ngp_ctx_set_scene(handle, _cabi.ptr(bf), self.cascade, self.grid_size, arr, float(self.min_near)),
                        "ngp_ctx_set_scene")
            ctx_owner.keys = dict(ctx_owner.keys or {}, scene=key)

    def run_cuda(self, rays_o, rays_d, dt_gamma=0, bg_color=None, perturb=False, force_all_rays=False, max_steps=1024,
                 fused=None, return_stats=False, return_uq=False, **kwargs):
        """nerf/renderer.py:261-386, inference branch.  rays_o, rays_d: [B, N, 3] -> dict(image [B,N,3], depth [B,N], ...)."""
        if self.training:
            raise NotImplementedError("run_cuda: the training branch (march_rays_train / composite_rays_train) is outside the "
                                      "inference hot path built here; call model.eval()")
        if self.bg_radius > 0:
            raise NotImplementedError("run_cuda: bg_radius > 0 (background network) is not on the validate.py path (bg_radius=-1)")
        _cabi.require_cuda(rays_o, rays_d)
        prefix = rays_o.shape[:-1]
        rays_o = rays_o.contiguous().view(-1, 3).float()
        rays_d = rays_d.contiguous().view(-1, 3).float()
        N = rays_o.shape[0]
        device = rays_o.device
        if bg_color is None:
            bg_color = 1
        fused = self.fused if fused is None else fused
        scalar_bg = not torch.is_tensor(bg_color)
        if torch.is_grad_enabled() and (rays_o.requires_grad or rays_d.requires_grad):
            # gradients w.r.t. the rays (Estimator.measurement_fn, nav/estimator_helpers.py:191-225): upstream only the sampler
            # path run() provides them (raymarching.py:161-230 has no backward); here the CUDA-ray render itself does
            if perturb or not scalar_bg or not self._configure_native(self._native):
                raise NotImplementedError("run_cuda: gradients w.r.t. rays need perturb=False, a scalar bg_color and the fusable NGP network")
            from .autograd import RenderRays
            image, depth, weights_sum = RenderRays.apply(rays_o, rays_d, self, float(dt_gamma), int(max_steps), float(bg_color))
            return {"depth": depth.view(*prefix), "image": image.view(*prefix, 3), "weights_sum": weights_sum, "sigmas": None, "rgbs": None}
        if (fused and 0 < N <= int(self.small_batch_rays) and not perturb and scalar_bg and not return_stats and not return_uq
                and self._configure_native(self._native)):
            return self._run_warp_per_ray(rays_o, rays_d, prefix, N, device, float(dt_gamma), int(max_steps), float(bg_color))
        if fused and not perturb and scalar_bg and self._configure_native(self._native):
            return self._run_fused(rays_o, rays_d, prefix, N, device, float(dt_gamma), int(max_steps), float(bg_color), return_stats,
                                   return_uq)
        return self._run_stepwise(rays_o, rays_d, prefix, N, device, dt_gamma, bg_color, perturb, max_steps)

    def _run_warp_per_ray(self, rays_o, rays_d, prefix, N, device, dt_gamma, max_steps, bg_color):
        """Small inference batches on the warp-per-ray kernel (see `small_batch_rays`): one launch, no workspace."""
        lib = _cabi.load()
        image = torch.empty(N, 3, dtype=torch.float32, device=device)
        depth = torch.empty(N, dtype=torch.float32, device=device)
        weights_sum = torch.empty(N, dtype=torch.float32, device=device)
        finals = torch.empty(N, 4, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            handle = self._native.get(device)
            self._set_scene(self._native, handle)
            _cabi.check(lib.ngp_render_diff_forward(handle, _cabi.ptr(rays_o), _cabi.ptr(rays_d), N, dt_gamma, max_steps, bg_color,
                                                    _cabi.ptr(image), _cabi.ptr(depth), _cabi.ptr(weights_sum), _cabi.ptr(finals), None,
                                                    _cabi.stream()), "ngp_render_diff_forward")
        return {"depth": depth.view(*prefix), "image": image.view(*prefix, 3), "weights_sum": weights_sum, "sigmas": None, "rgbs": None}

    def _run_fused(self, rays_o, rays_d, prefix, N, device, dt_gamma, max_steps, bg_color, return_stats, return_uq=False):
        lib = _cabi.load()
        image = torch.empty(N, 3, dtype=torch.float32, device=device)
        depth = torch.empty(N, dtype=torch.float32, device=device)
        weights_sum = torch.empty(N, dtype=torch.float32, device=device)
        spr = torch.empty(N, dtype=torch.int32, device=device) if return_stats else None
        stats = torch.zeros(4, dtype=torch.int64, device=device) if return_stats else None
        results = {"depth": depth.view(*prefix), "image": image.view(*prefix, 3), "weights_sum": weights_sum,
                   # the reference returns the LAST iteration's per-sample tensors (renderer.py:381-382); the fused loop only
                   # materialises them on request (return_uq=True)
                   "sigmas": None, "rgbs": None}
        with torch.cuda.device(device):   # the context, its workspace and the launch stream all belong to the tensors' device
            handle = self._native.get(device)
            self._set_scene(self._native, handle)
            if not return_uq:
                _cabi.check(lib.ngp_render(handle, _cabi.ptr(rays_o), _cabi.ptr(rays_d), N, dt_gamma, max_steps, bg_color,
                                           _cabi.ptr(image), _cabi.ptr(depth), _cabi.ptr(weights_sum), _cabi.ptr(spr), _cabi.ptr(stats),
                                           _cabi.stream()), "ngp_render")
            else:
                rows = C.c_uint64()
                _cabi.check(lib.ngp_render_sample_rows(handle, N, C.byref(rows)), "ngp_render_sample_rows")
                agg = torch.empty(N, dtype=torch.float32, device=device)
                dump = torch.empty(int(rows.value), 4, dtype=torch.float32, device=device)
                info = torch.empty(N, 4, dtype=torch.int32, device=device)
                hdr = torch.zeros(8, dtype=torch.int64, device=device)
                _cabi.check(lib.ngp_render_uq(handle, _cabi.ptr(rays_o), _cabi.ptr(rays_d), N, dt_gamma, max_steps, bg_color,
                                              _cabi.ptr(image), _cabi.ptr(depth), _cabi.ptr(weights_sum), _cabi.ptr(spr), _cabi.ptr(stats),
                                              _cabi.ptr(agg), _cabi.ptr(dump), int(rows.value), _cabi.ptr(info), _cabi.ptr(hdr),
                                              _cabi.stream()), "ngp_render_uq")
                results["aggregated_density"] = agg.view(*prefix)
                results["sigmas"], results["rgbs"] = self._last_iteration_samples(dump, info, hdr)
        if return_stats:
            results["samples_per_ray"] = spr
            results["stats"] = stats
        return results

    def _last_iteration_samples(self, dump, info, hdr):
        """The reference's `sigmas` / `rgbs` result keys (renderer.py:357-365, 381-382): the network outputs of the LAST iteration
        of its loop, row = (rank of the ray among the rays alive in that iteration, by ray id: the reference's compaction is
        stable) * n_step + k, padded to a multiple of 128 rows (march_rays' `align`); rows the marcher did not fill hold the
        network's output for xyz = dir = 0 (its buffers start as zeros).  Assembled from ngp_render_uq's per-sample dump of the
        last epoch (include/ngp_b200.h)."""
        tag, s, ran, kshift, slots, _, _, _ = (int(v) for v in hdr.cpu().tolist())
        device = dump.device
        zero = torch.zeros(1, 3, device=device)
        sigma0, rgb0 = self(zero, zero)
        sigma0 = float(self.density_scale) * sigma0.float()
        if ran <= 0:
            return None, None
        j = ran - 1                                                 # the last iteration the reference ran, within the last epoch
        K, R = 1 << kshift, 32 >> kshift
        rid = torch.nonzero(info[:, 0] == tag).flatten()            # ascending ray id = the reference's live-list order
        n, ne, t_death = (info[rid, c].long() for c in (1, 2, 3))
        alive = (ne >= j * s) & (t_death >= j) if j > 0 else torch.ones_like(n, dtype=torch.bool)   # survived iterations 0 .. j-1
        n, ne = n[alive], ne[alive]
        M = n.numel() * s
        padded = -(-M // 128) * 128 if M else 0
        sigmas = sigma0.expand(padded).clone()
        rgbs = rgb0.to(torch.float32).expand(padded, 3).clone()
        if M:
            i = j * s + torch.arange(s, device=device)[None, :]      # [rays, n_step] sample numbers within the epoch
            row = (n // R)[:, None] * (R * slots) + (i >> kshift) * 32 + (n % R)[:, None] * K + (i & (K - 1))
            valid = i < ne[:, None]
            out = dump[row.clamp_(0, dump.shape[0] - 1)]
            sigmas[:M] = torch.where(valid, out[..., 0], sigma0).reshape(-1)
            rgbs[:M] = torch.where(valid[..., None], out[..., 1:], rgb0.to(torch.float32)).reshape(-1, 3)
        return sigmas, rgbs

    def _run_stepwise(self, rays_o, rays_d, prefix, N, device, dt_gamma, bg_color, perturb, max_steps):
        with torch.no_grad():
            nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, self.aabb_train if self.training else self.aabb_infer,
                                                         self.min_near)
            weights_sum = torch.zeros(N, dtype=torch.float32, device=device)
            depth = torch.zeros(N, dtype=torch.float32, device=device)
            image = torch.zeros(N, 3, dtype=torch.float32, device=device)
            n_alive = N
            rays_alive = torch.arange(n_alive, dtype=torch.int32, device=device)
            rays_t = nears.clone()
            step = 0
            sigmas = rgbs = None
            while step < max_steps:
                n_alive = rays_alive.shape[0]
                if n_alive <= 0:
                    break
                n_step = max(min(N // n_alive, 8), 1)
                xyzs, dirs, deltas = raymarching.march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, self.bound,
                                                            self.density_bitfield, self.cascade, self.grid_size, nears, fars, 128,
                                                            perturb, dt_gamma, max_steps)
                sigmas, rgbs = self(xyzs, dirs)
                sigmas = self.density_scale * sigmas
                raymarching.composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image)
                rays_alive = raymarching.compact_alive(rays_alive)
                step += n_step
            image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
            depth = torch.clamp(depth - nears, min=0) / (fars - nears)
        return {"depth": depth.view(*prefix), "image": image.view(*prefix, 3), "sigmas": sigmas, "rgbs": rgbs,
                "weights_sum": weights_sum}

    def render(self, rays_o, rays_d, staged=False, max_ray_batch=4096, **kwargs):
        """nerf/renderer.py:549-588.  cuda_ray never stages: everything goes to run_cuda in one call.  cuda_ray=False runs the
        sampler `run()`, in chunks of max_ray_batch rays per batch element when staged."""
        if self.cuda_ray:
            return self.run_cuda(rays_o, rays_d, **kwargs)
        if not staged:
            return self.run(rays_o, rays_d, **kwargs)
        B, N = rays_o.shape[:2]
        device = rays_o.device
        depth = torch.empty((B, N), device=device)
        image = torch.empty((B, N, 3), device=device)
        aggregated_density = torch.empty((B, N), device=device)
        part = None
        for b in range(B):
            for head in range(0, N, max_ray_batch):
                tail = min(head + max_ray_batch, N)
                part = self.run(rays_o[b:b + 1, head:tail], rays_d[b:b + 1, head:tail], **kwargs)
                depth[b:b + 1, head:tail] = part["depth"]
                image[b:b + 1, head:tail] = part["image"]
                aggregated_density[b:b + 1, head:tail] = part["aggregated_density"]
        return {"depth": depth, "image": image, "rgbs": part["rgbs"], "sigmas": part["sigmas"], "aggregated_density": aggregated_density}

    @torch.no_grad()
    def render_poses(self, poses, intrinsics, H, W, dt_gamma=0, bg_color=1, max_steps=1024, perturb=False, return_stats=False, **kwargs):
        """Estimator.render_from_pose (nav/estimator_helpers.py:236-243) for a batch: poses [F,4,4] cam2world (CUDA, NGP frame),
        intrinsics (fx, fy, cx, cy) -> render() result with image [F, H*W, 3], depth [F, H*W].  On the fused path the rays
        (nerf/utils.py:52-116) are generated inside the render kernel (ngp_render_poses); otherwise by ngp_get_rays + render()."""
        scalar_bg = not torch.is_tensor(bg_color)
        if (self.cuda_ray and self.fused and not self.training and self.bg_radius <= 0 and not perturb and scalar_bg and not kwargs.get("return_uq")
                and kwargs.get("fused", True) and self._configure_native(self._native)):
            _cabi.require_cuda(poses)
            poses = poses.detach().to(torch.float32).contiguous()
            if tuple(poses.shape[1:]) != (4, 4):
                raise RuntimeError(f"render_poses: poses must be [F,4,4], got {tuple(poses.shape)}")
            F, device = poses.shape[0], poses.device
            N = F * H * W
            image = torch.empty(N, 3, dtype=torch.float32, device=device)
            depth = torch.empty(N, dtype=torch.float32, device=device)
            weights_sum = torch.empty(N, dtype=torch.float32, device=device)
            spr = torch.empty(N, dtype=torch.int32, device=device) if return_stats else None
            stats = torch.zeros(4, dtype=torch.int64, device=device) if return_stats else None
            fx, fy, cx, cy = (float(v) for v in intrinsics)
            with torch.cuda.device(device):
                handle = self._native.get(device)
                self._set_scene(self._native, handle)
                _cabi.check(_cabi.load().ngp_render_poses(handle, _cabi.ptr(poses), F, H, W, fx, fy, cx, cy, float(dt_gamma), int(max_steps),
                                                          float(bg_color), _cabi.ptr(image), _cabi.ptr(depth), _cabi.ptr(weights_sum),
                                                          _cabi.ptr(spr), _cabi.ptr(stats), _cabi.stream()), "ngp_render_poses")
            out = {"depth": depth.view(F, H * W), "image": image.view(F, H * W, 3), "weights_sum": weights_sum, "sigmas": None, "rgbs": None}
            if return_stats:
                out["samples_per_ray"], out["stats"] = spr, stats
            return out
        from .utils import get_rays
        r = get_rays(poses, intrinsics, H, W)
        return self.render(r["rays_o"], r["rays_d"], dt_gamma=dt_gamma, bg_color=bg_color, max_steps=max_steps, perturb=perturb,
                           return_stats=return_stats, **kwargs)

    @torch.no_grad()
    def render_poses_host(self, poses, intrinsics, H, W, image=None, depth=None, dt_gamma=0, bg_color=1, max_steps=1024):
        """The same through ONE C-ABI call with HOST buffers (ngp_render_poses_host): poses = CPU float32 tensor [F,4,4] (64 B per
        frame go up instead of 24 B per ray), image [F*H*W,3] / depth [F*H*W] = CPU float32 tensors (pinned for full PCIe speed;
        allocated if None).  Synchronous.  Returns (image, depth)."""
        if self.training or self.bg_radius > 0 or not self.cuda_ray:
            raise NotImplementedError("render_poses_host serves the cuda_ray inference path (model.eval(), bg_radius <= 0)")
        if not self._configure_native(self._native):
            raise RuntimeError("render_poses_host: this model cannot be served by the fused renderer")
        poses = poses.detach().to(torch.float32).contiguous()
        if poses.is_cuda or tuple(poses.shape[1:]) != (4, 4):
            raise RuntimeError("render_poses_host: poses must be a CPU tensor [F,4,4] (use render_poses for device poses)")
        F = poses.shape[0]
        n = F * H * W
        image = torch.empty(n, 3, dtype=torch.float32).pin_memory() if image is None else image
        depth = torch.empty(n, dtype=torch.float32).pin_memory() if depth is None else depth
        for t, shape in ((image, (n, 3)), (depth, (n,))):
            if t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or tuple(t.shape) != shape:
                raise RuntimeError(f"render_poses_host: output buffers must be contiguous CPU float32 {shape}")
        lib = _cabi.load()
        fx, fy, cx, cy = (float(v) for v in intrinsics)
        dev = self.density_bitfield.device
        with torch.cuda.device(dev):
            handle = self._native.get(dev)
            self._set_scene(self._native, handle)
            _cabi.check(lib.ngp_render_poses_host(handle, _cabi.ptr(poses), F, H, W, fx, fy, cx, cy, float(dt_gamma), int(max_steps),
                                                  float(bg_color), _cabi.ptr(image), _cabi.ptr(depth), _cabi.stream()), "ngp_render_poses_host")
        return image, depth

    @torch.no_grad()
    def mark_untrained_grid(self, poses, intrinsic, S=64):
        """nerf/renderer.py:388-449: density_grid = -1 for every cell that no training camera covers.  poses [B,4,4] cam2world,
        intrinsic (fx, fy, cx, cy).  One kernel over (cascade, Morton cell) that walks the poses in registers
        (ngp_mark_untrained_grid) instead of the reference's five nested Python loops; `S` (its block size) has no meaning here."""
        if not self.cuda_ray:
            return
        if isinstance(poses, np.ndarray):
            poses = torch.from_numpy(poses)
        dev = self.density_bitfield.device
        poses = poses.to(dev, torch.float32).contiguous()
        fx, fy, cx, cy = (float(v) for v in intrinsic)
        with torch.cuda.device(dev):
            _cabi.check(_cabi.load().ngp_mark_untrained_grid(_cabi.ptr(poses), poses.shape[0], fx, fy, cx, cy, self.grid_size, self.cascade,
                                                             float(self.bound), _cabi.ptr(self.density_grid), _cabi.stream()),
                        "ngp_mark_untrained_grid")

    def _cascade_sigmas(self, cas, coords, jitter):
        """density * density_scale at the jittered query points of cascade `cas` (renderer.py:484-494): all cells in Morton order
        (coords None) or the given cells."""
        dev = self.density_bitfield.device
        n = self.grid_size ** 3 if coords is None else coords.shape[0]
        xyzs = torch.empty(n, 3, dtype=torch.float32, device=dev)
        bound_c = float(min(2 ** cas, self.bound))
        _cabi.check(_cabi.load().ngp_occupancy_points(_cabi.ptr(coords), _cabi.ptr(jitter), n, self.grid_size, bound_c, _cabi.ptr(xyzs),
                                                      _cabi.stream()), "ngp_occupancy_points")
        sigmas = self.density(xyzs)["sigma"].reshape(-1).detach().float()
        return sigmas * self.density_scale

    @torch.no_grad()
    def update_extra_state(self, decay=0.95, S=128):
        """nerf/renderer.py:451-546.  The density grid lives in Morton order, so the update works in Morton order: per cascade ONE
        kernel writes the jittered query point of every cell at its Morton index (ngp_occupancy_points), ONE density query
        (ngp_density_forward) returns the cascade's row of the new grid -- no coordinate / index tensors, no scatter -- and the
        EMA-max is one in-place pass (ngp_density_grid_ema).  The random numbers are drawn exactly as the reference draws them (one
        torch.rand per S^3 block per cascade, in its loop order), so that a seeded call reproduces its grid."""
        if not self.cuda_ray:
            return
        dev = self.density_bitfield.device
        H, n_cas = self.grid_size, self.cascade
        lib = _cabi.load()
        with torch.cuda.device(dev):
            tmp_grid = torch.full_like(self.density_grid, -1.0)
            if self.iter_density < 16:                                           # full update (renderer.py:466-494)
                jitter = torch.empty(n_cas, H, H, H, 3, dtype=torch.float32, device=dev)
                cuts = [(h, min(h + S, H)) for h in range(0, H, S)]
                for x0, x1 in cuts:                                              # the reference's consumption order of the generator
                    for y0, y1 in cuts:
                        for z0, z1 in cuts:
                            for cas in range(n_cas):
                                jitter[cas, x0:x1, y0:y1, z0:z1] = torch.rand((x1 - x0) * (y1 - y0) * (z1 - z0), 3, device=dev).view(
                                    x1 - x0, y1 - y0, z1 - z0, 3)
                for cas in range(n_cas):
                    tmp_grid[cas] = self._cascade_sigmas(cas, None, jitter[cas].reshape(-1, 3))
            else:                                                                # partial update (renderer.py:497-523)
                N = H ** 3 // 4
                for cas in range(n_cas):
                    coords = torch.randint(0, H, (N, 3), device=dev)
                    occ = torch.nonzero(self.density_grid[cas] > 0).squeeze(-1)
                    occ = occ[torch.randint(0, occ.shape[0], [N], dtype=torch.long, device=dev)]   # duplicates allowed
                    coords = torch.cat([coords.int(), raymarching.morton3D_invert(occ)], dim=0).contiguous()
                    indices = torch.cat([raymarching.morton3D(coords[:N]).long(), occ], dim=0)
                    jitter = torch.rand(2 * N, 3, device=dev)
                    tmp_grid[cas, indices] = self._cascade_sigmas(cas, coords, jitter)
            _cabi.check(lib.ngp_density_grid_ema(_cabi.ptr(self.density_grid), _cabi.ptr(tmp_grid), tmp_grid.numel(), float(decay),
                                                 _cabi.stream()), "ngp_density_grid_ema")
            self.mean_density = torch.mean(self.density_grid.clamp(min=0)).item()   # -1 (never trained) counts as empty
            self.iter_density += 1
            density_thresh = min(self.mean_density, self.density_thresh)
            self.density_bitfield = raymarching.packbits(self.density_grid, density_thresh, self.density_bitfield)
        total_step = min(16, self.local_step)                                    # step-counter bookkeeping (renderer.py:540-544)
        if total_step > 0:
            self.mean_count = int(self.step_counter[:total_step, 0].sum().item() / total_step)
        self.local_step = 0
