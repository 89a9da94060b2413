"""CPU pins of tests/diff_oracle.py (the fp64 definition of the differentiable CUDA-ray render): its FORWARD against the C oracle
(oracle/ngp_oracle.c, itself pinned on the reference's golden vectors), its autograd DERIVATIVE against central differences of
itself.  tests/test_gpu_diff.py then holds the CUDA gradients to it."""
import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import synthetic as syn
from oracle import oracle as orc
from tests import diff_oracle as dorc
from tests.inputs import make_rays


def _rays_hitting(n, seed):
    """n camera rays of the cfg2-shaped orbit view that pass through occupied cells."""
    o, d = syn.pinhole_rays(syn.orbit_pose(), 64, 64)
    pick = np.random.default_rng(seed).permutation(o.shape[0])
    o, d = np.ascontiguousarray(o[pick], np.float32), np.ascontiguousarray(d[pick], np.float32)
    aabb = np.array([-1, -1, -1, 1, 1, 1], np.float32)
    nears, fars = orc.near_far_from_aabb(o, d, aabb)
    bitfield, _ = syn.stonehenge_bitfield()
    N = o.shape[0]
    _, _, deltas = orc.march_rays(N, 64, np.arange(N, dtype=np.int32), nears.copy(), o, d, 1.0, bitfield, 1, 128, nears, fars,
                                  dt_gamma=0.0, max_steps=1024)
    hits = (deltas[: N * 64, 0].reshape(N, 64) != 0).sum(1)
    keep = np.nonzero(hits >= 32)[0][:n]
    assert keep.size == n
    return torch.from_numpy(o[keep]), torch.from_numpy(d[keep]), nears[keep], fars[keep]


@pytest.mark.parametrize("variant", ["A", "B"])
def test_forward_matches_c_oracle(variant):
    w = syn.ngp_weights(0, variant=variant)
    net = dorc.Net(w)
    g = torch.Generator().manual_seed(1)
    x = torch.rand(2048, 3, generator=g) * 2 - 1
    x[:16] = 1.0
    x[16:32, 0] = 1.25                                   # outside the box: zero encoding
    d = torch.randn(2048, 3, generator=g)
    d = d / d.norm(dim=-1, keepdim=True)
    sig_ref, rgb_ref = orc.network_forward(x.numpy(), d.numpy(), w["table"], w["offsets"], w["S"], w["H"], w["w_sigma"], w["w_color"],
                                           w["num_layers"], w["num_layers_color"], bound=w["bound"], density_scale=1.0)
    sig, rgb = net.forward(x.double(), d.double())
    # the C oracle rounds activations to half between layers like the kernels; the fp64 restatement does not
    assert np.abs(sig.numpy() / sig_ref - 1).max() < 1e-2
    assert np.abs(rgb.numpy() - rgb_ref).max() < 2e-3


@pytest.mark.parametrize("variant", ["A", "B"])
def test_render_forward_and_derivative(variant):
    bitfield, _ = syn.stonehenge_bitfield()
    w = syn.ngp_weights(0, variant=variant)
    net = dorc.Net(w)
    N, S = 12, 1024
    o, d, nears, fars = _rays_hitting(N, 5)
    aabb = np.array([-1, -1, -1, 1, 1, 1], np.float32)
    alive = np.arange(N, dtype=np.int32)
    xyzs, dirs, deltas = orc.march_rays(N, S, alive, nears.copy(), o.numpy(), d.numpy(), w["bound"], bitfield, 1, 128, nears, fars,
                                        dt_gamma=0.0, max_steps=S)
    t, dt, count = dorc.samples_from_deltas(deltas, nears, N, S)
    assert int(count.max()) > 20
    S_used = int(count.max())
    t, dt = t[:, :S_used].contiguous(), dt[:, :S_used].contiguous()
    # positions recovered from t agree with the marcher's own xyz
    x_rec = (o.double()[:, None] + t[..., None] * d.double()[:, None]).clamp(-1, 1)
    x_m = torch.from_numpy(xyzs[: N * S]).view(N, S, 3)[:, :S_used].double()
    valid = torch.arange(S_used)[None] < count[:, None]
    assert float((x_rec - x_m).abs()[valid].max()) < 1e-5

    ref = orc.render(o.numpy(), d.numpy(), aabb, bitfield, w["table"], w["offsets"], w["S"], w["H"], w["w_sigma"], w["w_color"],
                     bound=w["bound"], max_steps=S, nl_sigma=w["num_layers"], nl_color=w["num_layers_color"],
                     density_scale=w["density_scale"], bg_color=1.0)
    od, dd = o.double().requires_grad_(True), d.double().requires_grad_(True)
    image, ws = dorc.render(net, od, dd, t, dt, count, bg=1.0)
    assert np.abs(image.detach().numpy() - ref["image"]).max() < 2e-3
    assert np.abs(ws.detach().numpy() - ref["weights_sum"]).max() < 2e-3

    G = torch.randn(N, 3, generator=torch.Generator().manual_seed(2)).double()
    (image * G).sum().backward()
    go, gd = od.grad.clone(), dd.grad.clone()
    assert float(go.abs().max()) > 0 and float(gd.abs().max()) > 0
    # central differences of the same fp64 function, ray by ray and axis by axis (rays are independent)
    eps = 2e-8
    with torch.no_grad():
        for which, base, grad in (("o", o.double(), go), ("d", d.double(), gd)):
            for ax in range(3):
                e = torch.zeros(N, 3, dtype=torch.float64)
                e[:, ax] = eps
                args_p = (base + e, d.double()) if which == "o" else (o.double(), base + e)
                args_m = (base - e, d.double()) if which == "o" else (o.double(), base - e)
                fp = (dorc.render(net, *args_p, t, dt, count)[0] * G).sum(1)
                fm = (dorc.render(net, *args_m, t, dt, count)[0] * G).sum(1)
                fd = (fp - fm) / (2 * eps)
                err = (fd - grad[:, ax]).abs()
                scale = grad.abs().max(1).values + 1e-9
                # a sample within eps of a grid-cell face or a unit within eps of its ReLU kink (variant A's pre-activations are ~1e-4)
                # makes that ray's difference quotient one-sided; the median over rays is robust against it
                assert float((err / scale).median()) < 1e-3, (which, ax, (err / scale).tolist())
