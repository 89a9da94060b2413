"""GPU: NeRFRenderer.run (cuda_ray=False, the hierarchical sampler `validate.py -O` selects) on the
device, against oracle/run_sampler.py -- the restatement pinned to the imported reference (tests/test_pyref_golden.py) --
driven by the SAME network: the oracle's sampler asks our model for density / colour, so what is compared is the sampler."""
import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import synthetic as syn
from oracle import run_sampler as rs

pytestmark = pytest.mark.gpu


class _OursAsNet:
    """Adapter: oracle/run_sampler.run's `net` interface on top of our CUDA model (CPU tensors in and out)."""

    def __init__(self, model, dev):
        self.m, self.dev = model, dev
        self.bound, self.min_near, self.density_scale = float(model.bound), float(model.min_near), float(model.density_scale)

    def density(self, x):
        with torch.no_grad():
            o = self.m.density(x.to(self.dev))
        self._geo_dtype = o["geo_feat"].dtype
        return o["sigma"].float().cpu(), o["geo_feat"].float().cpu()

    def rgb(self, d, geo, mask):
        # xyz is not used by the colour net (network_ff.py:92-134 takes it only for shapes)
        with torch.no_grad():
            c = self.m.color(torch.zeros_like(d).to(self.dev), d.to(self.dev), mask=mask.to(self.dev), geo_feat=geo.to(self.dev).to(self._geo_dtype))
        return c.float().cpu()


@pytest.mark.parametrize("variant", ["A", "B"])
def test_run_matches_the_pinned_sampler(cuda, scene, variant):
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant=variant)
    model = syn.build_model(w, bitfield, grid, device=cuda)
    model.cuda_ray = False                                       # what validate.py -O does (validate.py:115-118)
    o, d = syn.pinhole_rays(syn.orbit_pose(60.0), 24, 24)
    to, td = torch.from_numpy(o), torch.from_numpy(d)
    with torch.no_grad():
        out = model.render(to[None].to(cuda), td[None].to(cuda), staged=True, max_ray_batch=200, bg_color=1.0, perturb=False,
                           num_steps=64, upsample_steps=64)
    assert out["image"].shape == (1, 576, 3) and out["depth"].shape == (1, 576) and out["aggregated_density"].shape == (1, 576)
    ref = rs.render(_OursAsNet(model, cuda), to, td, max_ray_batch=200, num_steps=64, upsample_steps=64, bg_color=1.0)
    # tolerance: the two sides feed bit-identical sigmas through the same fp32 formulas on different devices (cumprod / exp)
    assert (out["image"][0].cpu() - ref["image"]).abs().max().item() <= 2e-4
    assert (out["depth"][0].cpu() - ref["depth"]).abs().max().item() <= 2e-4
    agg, ragg = out["aggregated_density"][0].cpu(), ref["aggregated_density"]
    assert ((agg - ragg).abs() / ragg.abs().clamp_min(1.0)).max().item() <= 1e-3
    un = model.render(to[None].to(cuda), td[None].to(cuda), staged=False, bg_color=1.0, perturb=False, num_steps=64, upsample_steps=64)
    assert (un["image"] - out["image"]).abs().max().item() <= 1e-5 and un["rgbs"].shape == (576, 128, 3)


def test_run_rejects_cpu_rays_and_training(cuda, scene):
    bitfield, grid = scene
    model = syn.build_model(syn.ngp_weights(0), bitfield, grid, device=cuda)
    model.cuda_ray = False
    o = torch.zeros(1, 4, 3)
    with pytest.raises(RuntimeError):
        model.render(o, o)
    model.train()
    with pytest.raises(NotImplementedError):
        model.render(o.to(cuda), o.to(cuda))
