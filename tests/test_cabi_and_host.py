"""CPU-only checks of the drop-in boundary and the host-side mirror of the reference interface (no compute calls):
the C-ABI library loads and exports every symbol include/ngp_b200.h declares, the Python packages keep the
reference's names / shapes / state-dict keys, and a CPU tensor is an error, never a fallback."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import _cabi, synthetic as syn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ngp_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ngp_[A-Za-z0-9_]+)\s*\(", src)))


def test_header_declares_what_the_binding_binds():
    assert set(declared_functions()) == set(_cabi.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_cabi.LIB_PATH):
        from nerfsafetyvalidation_b200.build import build
        build()
    lib = ctypes.CDLL(_cabi.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} is declared in include/ngp_b200.h but not exported"
    lib.ngp_abi_version.restype = ctypes.c_int
    assert lib.ngp_abi_version() == 1


def test_only_declared_symbols_are_exported():
    import subprocess
    if not os.path.exists(_cabi.LIB_PATH):
        pytest.skip("library not built")
    out = subprocess.run(["nm", "-D", "--defined-only", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    ours = {ln.split()[-1] for ln in out.splitlines() if " T " in ln and "ngp_" in ln.split()[-1]}
    assert ours == set(declared_functions())


def test_no_gpu_means_a_loud_error_not_a_fallback():
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU")
    lib = _cabi.load()
    h = ctypes.c_void_p()
    rc = lib.ngp_ctx_create(ctypes.byref(h))
    assert rc != 0 and "no CPU fallback" in _cabi.last_error()
    host = np.zeros((8, 3), np.int32)
    out = np.zeros(8, np.int32)
    rc = lib.ngp_morton3D(host.ctypes.data_as(ctypes.c_void_p), 8, out.ctypes.data_as(ctypes.c_void_p), None)
    assert rc != 0 and "device pointer" in _cabi.last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "nerfsafetyvalidation_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle/"
                assert "ngp_oracle" not in src and "/root/reference" not in src


def test_hashgrid_offsets_match_the_reference_layout():
    offs, pls = syn.hashgrid_offsets()
    assert abs(pls - 1.381912879967776) < 1e-12
    rows = np.diff(offs)
    assert rows[:5].tolist() == [4920, 13824, 32768, 85184, 216000] and (rows[5:] == 2 ** 19).all() and offs[-1] == 6119864
    offs22, _ = syn.hashgrid_offsets(log2_hashmap_size=22)
    assert offs22[-1] == 39625280


def test_module_surface_and_state_dict_names():
    from nerfsafetyvalidation_b200.nerf.network_ff import NeRFNetwork
    m = NeRFNetwork(encoding="hashgrid", bound=1, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10, bg_radius=-1)
    sd = m.state_dict()
    want = {"aabb_train": (6,), "aabb_infer": (6,), "density_grid": (1, 2097152), "density_bitfield": (262144,), "step_counter": (16, 2),
            "encoder.embeddings": (6119864, 2), "encoder.offsets": (17,), "sigma_net.weights": (7168,), "color_net.weights": (11264,)}
    for k, shape in want.items():
        assert k in sd and tuple(sd[k].shape) == shape, k
    assert m.cascade == 1 and m.grid_size == 128
    assert sd["encoder.embeddings"].dtype == torch.float32 and sd["density_bitfield"].dtype == torch.uint8
    import copy
    m2 = copy.deepcopy(m)          # NerfSimulator.py:37 deep-copies the model
    assert m2._native is not m._native
    m.eval()
    with pytest.raises(RuntimeError):
        m.render(torch.zeros(1, 4, 3), torch.zeros(1, 4, 3), bg_color=1, perturb=False)   # CPU tensors: error, not fallback
    m.train()
    with pytest.raises(NotImplementedError):
        m.render(torch.zeros(1, 4, 3), torch.zeros(1, 4, 3))


def test_package_names_of_the_reference_are_importable():
    from nerfsafetyvalidation_b200 import ffmlp, gridencoder, raymarching, shencoder
    for name in ("near_far_from_aabb", "march_rays", "composite_rays", "packbits", "morton3D", "morton3D_invert", "sph_from_ray",
                 "march_rays_train", "composite_rays_train"):
        assert callable(getattr(raymarching, name))
    assert hasattr(gridencoder, "GridEncoder") and hasattr(shencoder, "SHEncoder") and hasattr(ffmlp, "FFMLP")
    enc = gridencoder.GridEncoder(desired_resolution=2048)
    assert enc.output_dim == 32 and tuple(enc.embeddings.shape) == (6119864, 2)
    mlp = ffmlp.FFMLP(32, 16, 64, 2)
    assert mlp.weights.numel() == 7168


def test_synthetic_scene_is_deterministic():
    bf, grid = syn.stonehenge_bitfield()
    assert bf.shape == (262144,) and grid.shape == (1, 128 ** 3)
    frac = np.unpackbits(bf).mean()
    assert 0.05 < frac < 0.08
    import hashlib
    assert hashlib.sha1(bf.tobytes()).hexdigest() == hashlib.sha1(syn.stonehenge_bitfield()[0].tobytes()).hexdigest()
    o, d = syn.pinhole_rays(syn.orbit_pose(), 8, 8)
    assert np.allclose(np.linalg.norm(d, axis=-1), 1, atol=1e-6) and np.allclose(np.linalg.norm(o, axis=-1), np.hypot(1.6, 0.35), atol=1e-5)


def test_weight_caches_follow_writes_through_dot_data():
    """`param.data.copy_()` bumps neither _version nor data_ptr(); the fp16 casts the fused renderer reads must
    follow it anyway (content fingerprint in the cache key + invalidate_caches on load_state_dict / _apply)."""
    from nerfsafetyvalidation_b200.ffmlp import FFMLP
    from nerfsafetyvalidation_b200.gridencoder import GridEncoder
    enc = GridEncoder(num_levels=4, log2_hashmap_size=10)
    t0 = enc.half_table()
    assert enc.half_table() is t0                                   # unchanged parameter: the cast is reused
    enc.embeddings.data.copy_(torch.full_like(enc.embeddings, 0.25))
    t1 = enc.half_table()
    assert t1 is not t0 and float(t1.float().mean()) == 0.25
    mlp = FFMLP(32, 16, 64, 2)
    w0 = mlp.half_weights()
    sd = {k: torch.ones_like(v) for k, v in mlp.state_dict().items()}
    mlp.load_state_dict(sd)
    assert mlp._half is None                                        # dropped by _load_from_state_dict, not only by the fingerprint
    assert float(mlp.half_weights().float().min()) == 1.0 and mlp.half_weights() is not w0
    mlp.double()                                                    # _apply drops it as well
    assert mlp._half is None


def test_ctx_options_are_explicit_not_environment():
    """the library reads no environment variable; the knobs are per-context options."""
    src = open(os.path.join(ROOT, "nerfsafetyvalidation_b200", "csrc", "fused.cu")).read()
    assert "getenv" not in src
    from nerfsafetyvalidation_b200.nerf.renderer import _NativeCtx
    c = _NativeCtx()
    c.options["rows_per_ray"] = 64                                  # recorded without a context; applied when one is created
    assert c.handle is None and c.options == {"rows_per_ray": 64}


def test_raymarching_wrappers_refuse_cpu_tensors():
    """the reference moves CPU tensors with .cuda() (raymarching.py:34-35); here that is an error, as the header says."""
    from nerfsafetyvalidation_b200 import raymarching
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        raymarching.morton3D(torch.zeros(4, 3, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        raymarching.morton3D_invert(torch.zeros(4, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        raymarching.packbits(torch.zeros(1, 64), 0.5)
