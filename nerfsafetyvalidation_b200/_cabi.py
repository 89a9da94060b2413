"""ctypes binding of libngp_b200.so (include/ngp_b200.h).

The product has no CPU fallback: if the library is missing or a call fails, a RuntimeError is raised.
PyTorch is used by the callers only for device memory and streams; every kernel is ours.
"""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NGP_B200_LIB") or os.path.join(_PKG, "_lib", "libngp_b200.so")   # the override is for A/B builds only

_u32, _f32, _i32, _p, _u64 = C.c_uint32, C.c_float, C.c_int, C.c_void_p, C.c_uint64

# name -> argtypes ; every function returns int unless listed in _RESTYPES
_SIGNATURES = {
    "ngp_abi_version": [],
    "ngp_last_error": [],
    "ngp_launch_count": [],
    "ngp_near_far_from_aabb": [_p, _p, _p, _u32, _f32, _p, _p, _p],
    "ngp_get_rays": [_p, _u32, _u32, _u32, _f32, _f32, _f32, _f32, _p, _p, _p],
    "ngp_morton3D": [_p, _u32, _p, _p],
    "ngp_morton3D_invert": [_p, _u32, _p, _p],
    "ngp_packbits": [_p, _u32, _f32, _p, _p],
    "ngp_occupancy_points": [_p, _p, _u32, _u32, _f32, _p, _p],
    "ngp_density_grid_ema": [_p, _p, _u32, _f32, _p],
    "ngp_mark_untrained_grid": [_p, _u32, _f32, _f32, _f32, _f32, _u32, _u32, _f32, _p, _p],
    "ngp_march_rays": [_u32, _u32, _p, _p, _p, _p, _f32, _f32, _u32, _u32, _u32, _p, _p, _p, _p, _p, _p, _u32, _p],
    "ngp_composite_rays": [_u32, _u32, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "ngp_compact_scratch_bytes": [_u32],
    "ngp_compact_alive": [_p, _u32, _p, _p, _p, _p],
    "ngp_compact_alive_static": [_p, _u32, _p, _p, _p],
    "ngp_grid_encode_forward": [_p, _p, _i32, _p, _p, _u32, _u32, _u32, _u32, _f32, _u32, _u32, _i32, _i32, _i32, _p, _p],
    "ngp_grid_level_scales": [_u32, _f32, _u32, _p, _p],
    "ngp_sh_encode_forward": [_p, _p, _u32, _u32, _p],
    "ngp_ffmlp_inference": [_p, _p, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _p, _p],
    "ngp_ctx_create": [C.POINTER(_p)],
    "ngp_ctx_destroy": [_p],
    "ngp_ctx_set_network": [_p, _p, _p, _p, _u32, _f32, _u32, _p, _u32, _p, _u32, _f32, _f32, _p],
    "ngp_ctx_set_scene": [_p, _p, _u32, _u32, _p, _f32],
    "ngp_ctx_set_option": [_p, C.c_char_p, _u64],
    "ngp_ctx_enable_profile": [_p, _i32],
    "ngp_ctx_phase_cycles": [_p, _p, _i32],
    "ngp_network_forward": [_p, _p, _p, _u32, _i32, _p, _p, _p],
    "ngp_network_forward_tc": [_p, _p, _p, _u32, _i32, _p, _p, _p],
    "ngp_density_forward": [_p, _p, _u32, _i32, _p, _p, _p],
    "ngp_density_backward": [_p, _p, _p, _u32, _i32, _p, _p],
    "ngp_render_diff_forward": [_p, _p, _p, _u32, _f32, _u32, _f32, _p, _p, _p, _p, _p, _p],
    "ngp_render_diff_backward": [_p, _p, _p, _u32, _f32, _u32, _f32, _p, _p, _p, _p, _p, _p],
    "ngp_render": [_p, _p, _p, _u32, _f32, _u32, _f32, _p, _p, _p, _p, _p, _p],
    "ngp_render_uq": [_p, _p, _p, _u32, _f32, _u32, _f32, _p, _p, _p, _p, _p, _p, _p, _u64, _p, _p, _p],
    "ngp_render_sample_rows": [_p, _u32, C.POINTER(_u64)],
    "ngp_render_host": [_p, _p, _p, _u32, _f32, _u32, _f32, _p, _p, _p],
    "ngp_render_poses": [_p, _p, _u32, _u32, _u32, _f32, _f32, _f32, _f32, _f32, _u32, _f32, _p, _p, _p, _p, _p, _p],
    "ngp_render_poses_host": [_p, _p, _u32, _u32, _u32, _f32, _f32, _f32, _f32, _f32, _u32, _f32, _p, _p, _p],
}
_RESTYPES = {"ngp_last_error": C.c_char_p, "ngp_launch_count": _u64, "ngp_compact_scratch_bytes": _u64}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def load():
    """Loads the library (once). Raises RuntimeError if it has not been built -- never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if os.environ.get("NGP_B200_AUTOBUILD") == "1":
            from .build import build
            build()
        else:
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m nerfsafetyvalidation_b200.build` "
                "(or __graft_entry__.build()). nerfsafetyvalidation_b200 has no CPU / PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)
    if lib.ngp_abi_version() != 1:
        raise RuntimeError(f"libngp_b200.so ABI {lib.ngp_abi_version()} != 1 expected by the Python layer")
    _lib = lib
    return lib


def last_error() -> str:
    return load().ngp_last_error().decode("utf-8", "replace")


def check(rc: int, what: str = ""):
    if rc != 0:
        raise RuntimeError(f"{what or 'ngp_b200'}: {last_error()}")


def launch_count() -> int:
    return int(load().ngp_launch_count())


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("nerfsafetyvalidation_b200: expected a CUDA tensor (there is no CPU fallback)")


_FP_SAMPLES = 32


def fingerprint(*tensors):
    """Content fingerprint of parameter tensors: 32 evenly strided elements of each, read back in ONE device-to-host copy.
    Part of the key of the fp16 weight caches: `param.data.copy_(...)` changes neither `_version` nor `data_ptr()`, and a
    renderer that kept using the old cast would return wrong images without an error (the reference re-casts on every call)."""
    import torch
    parts = []
    for t in tensors:
        flat = t.detach().reshape(-1)
        n = flat.numel()
        if n == 0:
            continue
        step = max(1, n // _FP_SAMPLES)
        parts.append(flat[::step][:_FP_SAMPLES].float())
    if not parts:
        return b""
    return torch.cat(parts).cpu().numpy().tobytes()
