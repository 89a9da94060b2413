"""Drop-in for the reference's `raymarching` package (raymarching/raymarching.py), inference + setup functions.

Same names, argument order and return values as the reference wrappers; each call forwards to hand-written
sm_100a CUDA through the C ABI (include/ngp_b200.h).  Floating inputs are taken as fp32 exactly like the
reference's `custom_fwd(cast_inputs=torch.float32)` decorators.

Training-only entry points (`march_rays_train`, `composite_rays_train`) and `sph_from_ray` (only reached when
bg_radius > 0; validate.py uses -1) are outside this round's hot-path scope and raise.
"""
import torch

from .. import _cabi

__all__ = ["near_far_from_aabb", "sph_from_ray", "morton3D", "morton3D_invert", "packbits", "march_rays_train",
           "composite_rays_train", "march_rays", "composite_rays", "compact_alive"]


def _f32c(t):
    # the reference moves CPU tensors to the GPU silently (raymarching.py:34-35); here a CPU tensor is an error, as include/ngp_b200.h
    # and DESIGN.md say: there is no CPU path, and a hidden host-to-device copy per call is not one either
    _cabi.require_cuda(t)
    return t.float().contiguous()


def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
    """raymarching.py:19-49 -> (nears [N], fars [N]); misses get FLT_MAX in both."""
    lib = _cabi.load()
    rays_o = _f32c(rays_o).view(-1, 3)
    rays_d = _f32c(rays_d).view(-1, 3)
    aabb = _f32c(aabb)
    N = rays_o.shape[0]
    nears = torch.empty(N, dtype=torch.float32, device=rays_o.device)
    fars = torch.empty(N, dtype=torch.float32, device=rays_o.device)
    _cabi.check(lib.ngp_near_far_from_aabb(_cabi.ptr(rays_o), _cabi.ptr(rays_d), _cabi.ptr(aabb), N, float(min_near),
                                           _cabi.ptr(nears), _cabi.ptr(fars), _cabi.stream()), "near_far_from_aabb")
    return nears, fars


def sph_from_ray(rays_o, rays_d, radius):
    raise NotImplementedError("sph_from_ray is only reached with bg_radius > 0 (validate.py uses -1); out of this round's scope")


def morton3D(coords):
    """raymarching.py:83-104: coords int32 [N,3] in [0,1024) -> indices int32 [N]."""
    lib = _cabi.load()
    _cabi.require_cuda(coords)
    coords = coords.int().contiguous()
    N = coords.shape[0]
    indices = torch.empty(N, dtype=torch.int32, device=coords.device)
    _cabi.check(lib.ngp_morton3D(_cabi.ptr(coords), N, _cabi.ptr(indices), _cabi.stream()), "morton3D")
    return indices


def morton3D_invert(indices):
    """raymarching.py:106-126: indices int32 [N] -> coords int32 [N,3]."""
    lib = _cabi.load()
    _cabi.require_cuda(indices)
    indices = indices.int().contiguous()
    N = indices.shape[0]
    coords = torch.empty(N, 3, dtype=torch.int32, device=indices.device)
    _cabi.check(lib.ngp_morton3D_invert(_cabi.ptr(indices), N, _cabi.ptr(coords), _cabi.stream()), "morton3D_invert")
    return coords


def packbits(grid, thresh, bitfield=None):
    """raymarching.py:129-155: grid float [C, H^3] -> uint8 [C*H^3/8], bit i of a byte = grid[8n+i] > thresh."""
    lib = _cabi.load()
    grid = _f32c(grid)
    N = grid.shape[0] * grid.shape[1] // 8
    if bitfield is None:
        bitfield = torch.empty(N, dtype=torch.uint8, device=grid.device)
    _cabi.require_cuda(bitfield)
    _cabi.check(lib.ngp_packbits(_cabi.ptr(grid), N, float(thresh), _cabi.ptr(bitfield), _cabi.stream()), "packbits")
    return bitfield


def march_rays_train(*args, **kwargs):
    raise NotImplementedError("march_rays_train: training path, outside the inference hot path built here")


def composite_rays_train(*args, **kwargs):
    raise NotImplementedError("composite_rays_train: training path, outside the inference hot path built here")


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, density_bitfield, C, H, near, far,
               align=-1, perturb=False, dt_gamma=0, max_steps=1024):
    """raymarching.py:292-337 -> (xyzs [M,3], dirs [M,3], deltas [M,2]); M = n_alive*n_step padded UP to `align`
    (a full extra block when already a multiple, raymarching.py:326-327). Un-emitted rows stay zero."""
    lib = _cabi.load()
    rays_o = _f32c(rays_o).view(-1, 3)
    rays_d = _f32c(rays_d).view(-1, 3)
    rays_t, near, far = _f32c(rays_t), _f32c(near), _f32c(far)
    _cabi.require_cuda(rays_alive, density_bitfield)
    M = n_alive * n_step
    if align > 0:
        M += align - (M % align)
    dev = rays_o.device
    xyzs = torch.zeros(M, 3, dtype=torch.float32, device=dev)
    dirs = torch.zeros(M, 3, dtype=torch.float32, device=dev)
    deltas = torch.zeros(M, 2, dtype=torch.float32, device=dev)
    _cabi.check(lib.ngp_march_rays(int(n_alive), int(n_step), _cabi.ptr(rays_alive), _cabi.ptr(rays_t), _cabi.ptr(rays_o),
                                   _cabi.ptr(rays_d), float(bound), float(dt_gamma), int(max_steps), int(C), int(H),
                                   _cabi.ptr(density_bitfield), _cabi.ptr(near), _cabi.ptr(far), _cabi.ptr(xyzs),
                                   _cabi.ptr(dirs), _cabi.ptr(deltas), int(perturb), _cabi.stream()), "march_rays")
    return xyzs, dirs, deltas


def composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image):
    """raymarching.py:340-362: in place on rays_alive (dead -> -1), rays_t, weights_sum, depth, image."""
    lib = _cabi.load()
    sigmas, rgbs, deltas = _f32c(sigmas), _f32c(rgbs), _f32c(deltas)
    for t in (rays_alive, rays_t, weights_sum, depth, image):
        _cabi.require_cuda(t)
        if not t.is_contiguous():
            raise RuntimeError("composite_rays: in-place arguments must be contiguous")
    if rays_t.dtype != torch.float32 or weights_sum.dtype != torch.float32 or depth.dtype != torch.float32 or image.dtype != torch.float32:
        raise RuntimeError("composite_rays: rays_t / weights_sum / depth / image must be float32 (renderer.py:335-339)")
    _cabi.check(lib.ngp_composite_rays(int(n_alive), int(n_step), _cabi.ptr(rays_alive), _cabi.ptr(rays_t), _cabi.ptr(sigmas),
                                       _cabi.ptr(rgbs), _cabi.ptr(deltas), _cabi.ptr(weights_sum), _cabi.ptr(depth),
                                       _cabi.ptr(image), _cabi.stream()), "composite_rays")
    return tuple()


def compact_alive(rays_alive, n=None, return_count_tensor=False):
    """Stable `rays_alive[rays_alive >= 0]` (nerf/renderer.py:369) without the nonzero+gather round trip.

    Returns the compacted int32 tensor (one D2H read of the 4-byte count, which the reference loop needs anyway
    for `rays_alive.shape[0]`), or (padded_out, count_tensor) with no host sync if return_count_tensor."""
    lib = _cabi.load()
    _cabi.require_cuda(rays_alive)
    rays_alive = rays_alive.int().contiguous()
    if rays_alive.data_ptr() % 16:
        rays_alive = rays_alive.clone()  # the kernel reads int4 vectors
    n = rays_alive.shape[0] if n is None else int(n)
    out = torch.empty(max(n, 1), dtype=torch.int32, device=rays_alive.device)
    count = torch.zeros(1, dtype=torch.int32, device=rays_alive.device)
    scratch = torch.empty(int(lib.ngp_compact_scratch_bytes(max(n, 1))), dtype=torch.uint8, device=rays_alive.device)
    _cabi.check(lib.ngp_compact_alive(_cabi.ptr(rays_alive), n, _cabi.ptr(out), _cabi.ptr(count), _cabi.ptr(scratch),
                                      _cabi.stream()), "compact_alive")
    if return_count_tensor:
        return out, count
    return out[: int(count.item())]
