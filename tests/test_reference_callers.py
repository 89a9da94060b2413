"""The reference's untouched callers on our packages (north_star: "so validate.py's Monte Carlo and Cross Entropy Method stress tests
run it unchanged").

oracle/build_ref.py copies the reference's own `nerf/network_ff.py` and `validate.py` (unmodified, git-ignored build outputs) to
oracle/_ref/pyref.  Here INTEGRATION.md's `sys.modules` recipe is executed for real, then

  * the reference's OWN `NeRFNetwork` class (network_ff.py:11-148) is imported: every name it pulls in -- `encoding.get_encoder`,
    `activation.trunc_exp`, `ffmlp.FFMLP`, `.renderer.NeRFRenderer` -- resolves to this repository's modules; it loads our state dict
    (same keys / shapes) and renders through its own `forward` on our renderer;
  * the three closures validate.py builds at :288-291 (`density_fn`, `render_fn`, `get_rays_fn`) are executed from the file's own
    source text, with validate.py's own option defaults (`**vars(opt)` must be swallowed), against our model.
"""
import argparse
import importlib
import importlib.util
import os
import re
import sys
import types

import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import synthetic as syn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PYREF = os.path.join(ROOT, "oracle", "_ref", "pyref")
needs_pyref = pytest.mark.skipif(not os.path.exists(os.path.join(PYREF, "validate.py")), reason="oracle/_ref/pyref not built (python oracle/build_ref.py)")

RECIPE_NAMES = ("raymarching", "gridencoder", "shencoder", "ffmlp", "encoding", "activation")


@pytest.fixture
def swapped_modules():
    """INTEGRATION.md section 1, verbatim in effect; undone afterwards so other tests see a clean sys.modules."""
    touched = list(RECIPE_NAMES) + ["nerf", "nerf.network_ff", "nerf.renderer", "nerf.utils"]
    saved = {k: sys.modules.get(k) for k in touched}
    for name in RECIPE_NAMES:
        sys.modules[name] = importlib.import_module(f"nerfsafetyvalidation_b200.{name}")
    sys.modules["nerf.renderer"] = importlib.import_module("nerfsafetyvalidation_b200.nerf.renderer")
    sys.modules["nerf.utils"] = importlib.import_module("nerfsafetyvalidation_b200.nerf.utils")
    pkg = types.ModuleType("nerf")
    pkg.__path__ = [os.path.join(PYREF, "nerf")]
    sys.modules["nerf"] = pkg
    sys.modules.pop("nerf.network_ff", None)
    yield
    for k, v in saved.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v


def reference_network_class():
    spec = importlib.util.spec_from_file_location("nerf.network_ff", os.path.join(PYREF, "nerf", "network_ff.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["nerf.network_ff"] = mod
    spec.loader.exec_module(mod)
    return mod.NeRFNetwork


def validate_py_options():
    """validate.py's own argparse block (:59-108), executed from its source text, parsed with the flags the stress tests use."""
    src = open(os.path.join(PYREF, "validate.py")).read()
    block = re.search(r"^(\s*)parser = argparse\.ArgumentParser\(\)\n(.*?)\n\s*opt = parser\.parse_args\(\)", src, re.S | re.M)
    assert block, "validate.py's option block was not found"
    lines = ["parser = argparse.ArgumentParser()"] + [ln.strip() for ln in block.group(2).splitlines() if ln.strip().startswith("parser.add_argument")]
    ns = {"argparse": argparse, "random": __import__("random")}
    exec("\n".join(lines), ns)
    return ns["parser"].parse_args(["scene_dir", "--ff", "--cuda_ray", "--bound", "1", "--dt_gamma", "0"])


def validate_py_closures():
    src = open(os.path.join(PYREF, "validate.py")).read()
    out = {}
    for name in ("density_fn", "render_fn", "get_rays_fn"):
        m = re.search(rf"^\s*({name} = lambda .*)$", src, re.M)
        assert m, f"validate.py no longer defines {name}"
        out[name] = m.group(1)
    return out


@needs_pyref
def test_recipe_resolves_every_import_of_the_reference_network(swapped_modules):
    """CPU: the reference's network_ff.py imports cleanly on our modules and builds the same parameter set."""
    Net = reference_network_class()
    from nerfsafetyvalidation_b200.nerf import network_ff as ours
    assert Net is not ours.NeRFNetwork and Net.__mro__[1] is sys.modules["nerf.renderer"].NeRFRenderer
    ref_model = Net(encoding="hashgrid", bound=1, cuda_ray=True, density_scale=1)
    our_model = ours.NeRFNetwork(encoding="hashgrid", bound=1, cuda_ray=True, density_scale=1)
    a, b = ref_model.state_dict(), our_model.state_dict()
    assert {k: tuple(v.shape) for k, v in a.items()} == {k: tuple(v.shape) for k, v in b.items()}
    opt = validate_py_options()
    assert opt.cuda_ray and opt.ff and opt.max_steps == 1024
    assert set(validate_py_closures()) == {"density_fn", "render_fn", "get_rays_fn"}


@needs_pyref
@pytest.mark.gpu
def test_reference_network_and_validate_closures_run_on_our_packages(swapped_modules, cuda, scene):
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant="B")
    ours = syn.build_model(w, bitfield, grid, device=cuda)
    Net = reference_network_class()
    opt = validate_py_options()
    # validate.py:143-156 constructs the model with these keywords
    model = Net(encoding="hashgrid", bound=opt.bound, cuda_ray=opt.cuda_ray, density_scale=w["density_scale"], min_near=opt.min_near,
                density_thresh=opt.density_thresh, bg_radius=opt.bg_radius)
    missing = model.load_state_dict(ours.state_dict(), strict=False)      # nerf/utils.py:1017 loads checkpoints with strict=False
    assert not missing.missing_keys and not missing.unexpected_keys
    model = model.to(cuda).eval()

    H = W = 48
    dataset = types.SimpleNamespace(intrinsics=np.array([60.0, 60.0, W / 2, H / 2]), H=H, W=W)
    device = cuda
    pose = torch.from_numpy(syn.orbit_pose()[None].astype(np.float32)).to(cuda)
    ns = {"model": model, "opt": opt, "dataset": dataset, "torch": torch, "device": device,
          "get_rays": sys.modules["nerf.utils"].get_rays,
          "rot": torch.tensor([[0., 0., 1.], [1., 0., 0.], [0., 1., 0.]], device=cuda, dtype=torch.float32)}
    for line in validate_py_closures().values():
        exec(line, ns)
    with torch.no_grad():
        rays = ns["get_rays_fn"](pose)
        out = ns["render_fn"](rays["rays_o"], rays["rays_d"])            # the reference's forward() on our encoder / FFMLP / renderer
        x = torch.rand(1000, 3, device=cuda) * 1.6 - 0.8
        sigma = ns["density_fn"](x)
        ref = ours.render(rays["rays_o"], rays["rays_d"], bg_color=1., perturb=False, dt_gamma=opt.dt_gamma, max_steps=opt.max_steps)
        sigma_ours = ours.density(x.reshape(-1, 3) @ ns["rot"])["sigma"].reshape(x.shape[:-1])
    assert out["image"].shape == (1, H * W, 3) and out["depth"].shape == (1, H * W)
    # their Python forward composes our stand-alone ops (fp16 between them); ours is the fused kernel: both within the MLP tolerance
    assert float((out["image"] - ref["image"]).abs().max()) <= 4e-3
    hit = ~torch.isnan(ref["depth"])
    assert torch.equal(torch.isnan(out["depth"]), ~hit) and float((out["depth"] - ref["depth"])[hit].abs().max()) <= 4e-3
    assert float(((sigma - sigma_ours).abs() / sigma_ours.abs().clamp_min(1e-6)).max()) <= 2e-2
