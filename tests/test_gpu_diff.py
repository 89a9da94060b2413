"""GPU tests of the differentiable small-batch path (include/ngp_b200.h: ngp_render_diff_forward /
ngp_render_diff_backward / ngp_density_backward), through the reference-facing surface: `model.render(rays_o, rays_d, ...)` with
rays that require grad, and `model.density(x)` with points that require grad.

Oracles:
  * forward: the inference renderer (already held to the reference's extensions by test_gpu_parity.py / test_gpu_configs.py);
  * derivative: tests/diff_oracle.py, the fp64 autograd definition (pinned on the CPU by tests/test_diff_oracle.py);
  * the reference's OWN backward kernels (grid_encode_backward + dy_dx, sh_encode_backward, ffmlp_backward, composite_rays_train_backward)
    driven by its OWN Python (tests/refext.reference_training_stack()), with the one line the reference lacks, x = o + t d, in torch.
Tolerances are relative L2 errors of the gradient over the batch; they are written next to each assert with what was measured.
"""
import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import raymarching as our_rm
from nerfsafetyvalidation_b200 import synthetic as syn
from tests import diff_oracle as dorc
from tests import refext

pytestmark = pytest.mark.gpu

N_RAYS, MAX_STEPS = 384, 1024


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300))


def all_active_weights():
    """Variant "C": a network with NO ReLU kinks on the data -- non-negative table and weights, so every hidden unit is active at
    every sample.  The derivative of a ReLU network is piecewise constant: wherever the kernels' fp16-rounded forward and the unrounded
    fp64 definition disagree on the sign of one pre-activation (about 1 unit in 1000 with variants A / B), that sample's gradient
    differs by ~1/sqrt(active units).  Variant C removes that effect and leaves the arithmetic of the backward chain itself
    (compositor, sigmoid / exp, both MLPs, SH Jacobian, trilinear input gradient, x = o + t d) to be compared tightly."""
    w = dict(syn.ngp_weights(0, variant="B"))
    w["density_scale"] = 1.0
    table = np.abs(w["table_f32"]).astype(np.float16)
    ws = np.abs(w["w_sigma_f32"]).astype(np.float32)
    ws[2048:-1024] *= 0.1                               # hidden activations stay O(1) ...
    ws[-1024:] *= 0.2                                   # ... and h0 ~ 1.7: sigma ~ 5, weights_sum ~ 0.75 after ~80 samples
    wc = np.abs(w["w_color_f32"]).astype(np.float32)
    wc[2048:] *= 0.1                                    # keeps the pre-sigmoid output O(1)
    ws, wc = ws.astype(np.float16), wc.astype(np.float16)
    w.update(table=table, table_f32=table.astype(np.float32), w_sigma=ws, w_sigma_f32=ws.astype(np.float32), w_color=wc,
             w_color_f32=wc.astype(np.float32))
    return w


@pytest.fixture(scope="module")
def world(cuda, scene):
    bitfield, grid = scene
    o, d = syn.pinhole_rays(syn.orbit_pose(), 96, 96)
    pick = np.random.default_rng(11).permutation(o.shape[0])[:N_RAYS]
    o, d = np.ascontiguousarray(o[pick], np.float32), np.ascontiguousarray(d[pick], np.float32)
    o[-16:], d[-16:] = 3.0, np.float32(1 / np.sqrt(3))          # rays that miss the box: background pixels, NaN depth, zero gradient
    o[-32:-16] = np.float32([0.0, 0.9, 0.0])                    # rays from inside the box through empty space only
    d[-32:-16] = np.float32([0.0, 1.0, 0.0])
    out = dict(o=torch.from_numpy(o).to(cuda), d=torch.from_numpy(d).to(cuda), bitfield=bitfield)
    for var in "ABC":
        w = all_active_weights() if var == "C" else syn.ngp_weights(0, variant=var)
        out["w" + var] = w
        out["m" + var] = syn.build_model(w, bitfield, grid, device=cuda)
    out["G"] = torch.randn(N_RAYS, 3, generator=torch.Generator().manual_seed(3)).to(cuda)
    return out


def our_grads(model, o, d, G, g_ws=None, **kw):
    o = o.clone().requires_grad_(True)
    d = d.clone().requires_grad_(True)
    res = model.render(o[None], d[None], staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS, **kw)
    loss = (res["image"][0] * G).sum()
    if g_ws is not None:
        loss = loss + (res["weights_sum"] * g_ws).sum()
    loss.backward()
    return res, o.grad, d.grad


def marched_samples(model, o, d):
    """(t, dt, count) float64 [N,S] of every ray's samples, from the stand-alone marcher (bit-exact with the reference's)."""
    N = o.shape[0]
    nears, fars = our_rm.near_far_from_aabb(o, d, model.aabb_infer, model.min_near)
    alive = torch.arange(N, dtype=torch.int32, device=o.device)
    _, _, deltas = our_rm.march_rays(N, MAX_STEPS, alive, nears.clone(), o, d, model.bound, model.density_bitfield, model.cascade,
                                     model.grid_size, nears, fars, -1, False, 0, MAX_STEPS)
    t, dt, count = dorc.samples_from_deltas(deltas, nears, N, MAX_STEPS)
    S = max(int(count.max()), 1)
    return t[:, :S].contiguous(), dt[:, :S].contiguous(), count


@pytest.mark.parametrize("var", ["A", "B", "C"])
def test_forward_equals_inference_render(world, var):
    model, o, d = world["m" + var], world["o"], world["d"]
    with torch.no_grad():
        ref = model.render(o[None], d[None], staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS)
    res, _, _ = our_grads(model, o, d, world["G"])
    assert res["image"].requires_grad and not res["depth"].requires_grad
    assert_same_render(res, ref)


def assert_same_render(res, ref):
    """Same march / network / compositor functions as the inference renderer (mma.sync instead of tcgen05 accumulation order: measured
    <= 3e-6 on A / C, <= 6e-5 on B), except that a ray is marched as ONE chunk instead of the reference's batch-dependent chunks of
    n_step samples: where the reference's re-summed t is an ulp off after a large empty gap, one sample can land on the other side of
    a cell face.  Measured: 2 of 65 536 rays, 1.4e-3 on those.  So: nearly every ray within 1e-4, every ray within 5e-3."""
    for key in ("image", "weights_sum"):
        diff = (res[key] - ref[key]).abs()
        assert float(diff.max()) < 5e-3, key
        assert float((diff > 1e-4).float().mean()) < 1e-3, key
    dm = (res["depth"] - ref["depth"]).abs()
    assert torch.equal(torch.isnan(res["depth"]), torch.isnan(ref["depth"]))
    assert float((dm[~torch.isnan(dm)] > 1e-4).float().mean()) < 1e-3


def test_small_batch_route(world):
    """Opt-in: inference batches of at most `small_batch_rays` rays take the warp-per-ray kernel (one launch instead of the persistent
    kernel's cooperative launch + grid barriers); larger batches and calls that want per-ray statistics stay on the persistent kernel."""
    from nerfsafetyvalidation_b200 import _cabi
    model, o, d = world["mB"], world["o"], world["d"]
    kw = dict(staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS)
    with torch.no_grad():
        ref = model.render(o[None], d[None], **kw)
        try:
            model.small_batch_rays = 4096
            before = _cabi.launch_count()
            res = model.render(o[None], d[None], **kw)
            assert _cabi.launch_count() == before + 1
            stats = model.render(o[None], d[None], return_stats=True, **kw)           # persistent kernel: it owns the statistics
            assert stats["samples_per_ray"] is not None and torch.equal(stats["image"], ref["image"])
            model.small_batch_rays = 16
            assert torch.equal(model.render(o[None], d[None], **kw)["image"], ref["image"])   # above the limit: persistent kernel
        finally:
            model.small_batch_rays = 0
    assert_same_render(res, ref)
    assert not res["image"].requires_grad


# aggregate relative L2 error of the batch's gradient / median over rays of the per-ray relative error
# / |median over rays of (g . g_def) / (g_def . g_def) - 1| (a systematic scale error would show here)
RAY_TOL = {"A": (1e-1, 5e-2, 2e-2), "B": (1e-1, 3e-2, 5e-3), "C": (5e-2, 5e-3, 5e-3)}


def per_row_rel(a, b):
    return (a.double() - b.double()).norm(dim=-1) / b.double().norm(dim=-1).clamp_min(1e-300)


@pytest.mark.parametrize("var", ["A", "B", "C"])
def test_ray_gradients_match_fp64_definition(world, var, cuda):
    model, w, o, d, G = world["m" + var], world["w" + var], world["o"], world["d"], world["G"]
    g_ws = torch.linspace(-1, 1, N_RAYS, device=cuda)
    _, go, gd = our_grads(model, o, d, G, g_ws)
    t, dt, count = marched_samples(model, o, d)
    assert int((count > 0).sum()) > N_RAYS // 4
    net = dorc.Net(w, device=cuda)
    od, dd = o.double().requires_grad_(True), d.double().requires_grad_(True)
    image, ws = dorc.render(net, od, dd, t, dt, count, bg=1.0)
    ((image * G.double()).sum() + (ws * g_ws.double()).sum()).backward()
    missed = count == 0
    assert int(missed.sum()) >= 16
    assert float(go[missed].abs().max()) == 0.0 and float(gd[missed].abs().max()) == 0.0
    eo, ed = rel_l2(go, od.grad), rel_l2(gd, dd.grad)
    hit = ~missed
    mo, md = float(per_row_rel(go, od.grad)[hit].median()), float(per_row_rel(gd, dd.grad)[hit].median())
    print(f"variant {var}: rel L2 error d/d rays_o {eo:.3e}, d/d rays_d {ed:.3e}; per-ray median {mo:.3e} / {md:.3e}; |g_o| {float(od.grad.norm()):.3e}")
    ratio = float(((go.double() * od.grad).sum(-1) / (od.grad * od.grad).sum(-1).clamp_min(1e-300))[hit].median())
    print(f"variant {var}: median projection ratio {ratio:.6f}")
    # measured (B200), aggregate / per-ray median: A 4.0e-2 / 3.2e-2, B 4.0e-2 / 6e-3 (ReLU sign disagreements between the fp16-rounded
    # forward and the fp64 definition, see all_active_weights; the reference's own backward kernels sit at 6e-2 from the same
    # definition), C 3.0e-2 / 1e-3 (what is left in C's aggregate: samples that the fp32 march puts on the other side of a grid-cell
    # face than the fp64 restatement of t -- the trilinear derivative jumps there); projection ratio 1.00005 (B), 0.99998 (C)
    tol, tol_med, tol_ratio = RAY_TOL[var]
    assert eo < tol and ed < tol
    assert mo < tol_med and md < tol_med
    assert abs(ratio - 1) < tol_ratio


def test_gradient_is_independent_of_the_loss_scale(world):
    """Every fp32 -> fp16 conversion of the backward is preceded by a per-row power-of-two renormalisation: scaling the upstream
    gradient by 2^-24 (a pixel loss's magnitude, where a plain half backward underflows) scales the result by exactly 2^-24."""
    model, o, d, G = world["mB"], world["o"], world["d"], world["G"]
    _, go, gd = our_grads(model, o, d, G)
    _, go_s, gd_s = our_grads(model, o, d, G * 2.0 ** -24)
    assert float(go.abs().max()) > 0
    assert torch.equal(go_s * 2.0 ** 24, go) and torch.equal(gd_s * 2.0 ** 24, gd)


@pytest.mark.parametrize("var", ["A", "B", "C"])
def test_density_gradient(world, var, cuda):
    model, w = world["m" + var], world["w" + var]
    g = torch.Generator().manual_seed(8)
    x = (torch.rand(5000, 3, generator=g) * 2 - 1)
    x[:8] = 1.0
    x[8:16, 1] = 1.5                                          # outside: sigma = exp(h(0)), gradient 0
    x = x.to(cuda)
    gs = torch.randn(5000, generator=g).to(cuda)
    xr = x.clone().requires_grad_(True)
    out = model.density(xr)
    assert out["sigma"].requires_grad and not out["geo_feat"].requires_grad
    with torch.no_grad():
        plain = model.density(x)
    assert torch.equal(out["sigma"].detach(), plain["sigma"]) and torch.equal(out["geo_feat"], plain["geo_feat"])
    (out["sigma"] * gs).sum().backward()
    net = dorc.Net(w, device=cuda)
    xd = x.double().requires_grad_(True)
    sig, _ = net.density(xd)
    (sig * gs.double()).sum().backward()
    assert float(xr.grad[8:16].abs().max()) == 0.0
    e = rel_l2(xr.grad, xd.grad)
    med = float(per_row_rel(xr.grad[16:], xd.grad[16:]).median())
    print(f"variant {var}: density gradient rel L2 error {e:.3e}, per-point median {med:.3e}")
    # measured (B200): A 4.1e-2, B 1.6e-2 (ReLU sign disagreements on a few points, see all_active_weights), C 4.1e-3; medians 3e-4 .. 4e-4
    assert e < (2e-2 if var == "C" else 1e-1)
    assert med < 5e-3


needs_train_stack = pytest.mark.skipif(
    not all(refext.available(n) for n in ("_raymarching", "_gridencoder", "_shencoder", "_ffmlp")),
    reason="oracle/_ref extensions not built (python oracle/build_ref.py)")


def build_reference_model(w, ours, device):
    """The reference's NeRFNetwork (its own network_ff.py / renderer.py / gridencoder / shencoder / ffmlp Python on the rebuilt extensions)
    carrying our model's state dict, in train mode (ffmlp.py:162: only the training kernels keep what the backward needs)."""
    stack = refext.reference_training_stack()
    assert stack is not None, "oracle/_ref/pyref is incomplete (python oracle/build_ref.py)"
    Net, rm = stack
    ref = Net(encoding="hashgrid", bound=w["bound"], cuda_ray=True, density_scale=w["density_scale"], min_near=0.2, density_thresh=10.0,
              bg_radius=-1, num_layers=w["num_layers"], num_layers_color=w["num_layers_color"])
    missing = ref.load_state_dict(ours.state_dict(), strict=False)
    assert not missing.missing_keys and not missing.unexpected_keys
    return ref.to(device).train(), rm


def reference_parts_grads(ref, rm, o, d, G, g_ws):
    """The reference's training-mode CUDA-ray render (nerf/renderer.py:285-327: march_rays_train -> NeRFNetwork.forward under autocast
    -> composite_rays_train) through its OWN Python and backward kernels, with the sample positions rebuilt as x = o + t d in torch so
    that autograd reaches the rays (march_rays_train returns xyzs without a backward, raymarching.py:161-230)."""
    device = o.device
    nears, fars = rm.near_far_from_aabb(o, d, ref.aabb_infer, ref.min_near)
    xyzs, dirs, deltas, rays = rm.march_rays_train(o, d, ref.bound, ref.density_bitfield, ref.cascade, ref.grid_size, nears, fars, None, -1,
                                                   False, 128, True, 0, MAX_STEPS)
    M = xyzs.shape[0]
    order = torch.argsort(rays[:, 1])
    idx, cnt = rays[order, 0].long(), rays[order, 2].long()
    total = int(cnt.sum())
    rid = torch.zeros(M, dtype=torch.long, device=device)
    rid[:total] = torch.repeat_interleave(idx, cnt)
    live = torch.arange(M, device=device) < total
    t = (((xyzs - o[rid]) * d[rid]).sum(-1) / (d[rid] * d[rid]).sum(-1)) * live         # the parameter each sample was taken at
    orq, drq = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
    x = orq[rid] + t[:, None] * drq[rid]
    assert float((x.detach() - xyzs)[live].abs().max()) < 1e-5
    with torch.autocast("cuda", dtype=torch.float16):
        sigmas, rgbs = ref(x, drq[rid])
        sigmas = ref.density_scale * sigmas
        ws, depth, image = rm.composite_rays_train(sigmas, rgbs, deltas, rays)
        image = image + (1 - ws).unsqueeze(-1) * 1.0
    ((image * G).sum() + (ws * g_ws).sum()).backward()
    return image.detach(), orq.grad, drq.grad


@needs_train_stack
@pytest.mark.parametrize("var", ["A", "B", "C"])
def test_ray_gradients_against_the_reference_backward_kernels(world, var, cuda):
    model, w, o, d, G = world["m" + var], world["w" + var], world["o"], world["d"], world["G"]
    g_ws = torch.linspace(-1, 1, N_RAYS, device=cuda)
    res, go, gd = our_grads(model, o, d, G, g_ws)
    ref, rm = build_reference_model(w, model, cuda)
    img_ref, go_ref, gd_ref = reference_parts_grads(ref, rm, o, d, G, g_ws)
    assert float((res["image"][0] - img_ref).abs().max()) < 1e-3          # north_star's image tolerance
    # the fp64 definition, to say which side the difference comes from
    t, dt, count = marched_samples(model, o, d)
    net = dorc.Net(w, device=cuda)
    od, dd = o.double().requires_grad_(True), d.double().requires_grad_(True)
    image, ws = dorc.render(net, od, dd, t, dt, count, bg=1.0)
    ((image * G.double()).sum() + (ws * g_ws.double()).sum()).backward()
    ours_ref = (rel_l2(go, go_ref), rel_l2(gd, gd_ref))
    ours_def = (rel_l2(go, od.grad), rel_l2(gd, dd.grad))
    ref_def = (rel_l2(go_ref, od.grad), rel_l2(gd_ref, dd.grad))
    print(f"variant {var}: ours vs reference kernels {ours_ref[0]:.3e} / {ours_ref[1]:.3e}; ours vs fp64 {ours_def[0]:.3e} / {ours_def[1]:.3e}; "
          f"reference kernels vs fp64 {ref_def[0]:.3e} / {ref_def[1]:.3e}")
    # the reference's backward is fp16 throughout (half dy_dx, half backward buffers) and its forward rounds differently from ours, so
    # with A / B both sides disagree with the fp64 definition -- and with each other -- on a few ReLU signs: measured (B200) ours vs
    # reference 6.6e-2, reference vs fp64 6.2e-2, ours vs fp64 4.0e-2.  Variant C (no kinks) compares the arithmetic alone: ours vs
    # the reference's kernels 5.5e-3 / 4.9e-3, while BOTH sit at 3.0e-2 from the fp64 definition (its fp64 restatement of t puts a few
    # samples on the other side of a grid-cell face than the fp32 x both kernel sets use).
    assert max(ours_ref) < (2e-2 if var == "C" else 1.5e-1)
    assert max(ours_def) <= max(ref_def) * 1.5 + 1e-3


@needs_train_stack
@pytest.mark.parametrize("var", ["A", "B", "C"])
def test_density_gradient_against_the_reference_backward_kernels(world, var, cuda):
    """NeRFNetwork.density under autograd, reference side: gridencoder/grid.py:62-90 (dy_dx) + ffmlp.py:51-96 on the rebuilt kernels."""
    model, w = world["m" + var], world["w" + var]
    ref, _ = build_reference_model(w, model, cuda)
    g = torch.Generator().manual_seed(9)
    x = (torch.rand(4096, 3, generator=g) * 2 - 1).to(cuda)
    gs = torch.randn(4096, generator=g).to(cuda)
    xr = x.clone().requires_grad_(True)
    (model.density(xr)["sigma"] * gs).sum().backward()
    xq = x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.float16):
        sig = ref.density(xq)["sigma"]
    (sig.float() * gs).sum().backward()
    e = rel_l2(xr.grad, xq.grad)
    med = float(per_row_rel(xr.grad, xq.grad).median())
    print(f"variant {var}: density gradient, ours vs reference kernels {e:.3e}, per-point median {med:.3e}")
    # measured (B200): A 7.9e-2, B 5.2e-2 (ReLU sign disagreements on a few points + the reference's half dy_dx), C 9.8e-4
    assert e < (5e-3 if var == "C" else 1.5e-1)
    assert med < 2e-2


def test_two_cascades_and_cone_stepping(cuda):
    """bound = 2 (two occupancy cascades, mip levels in the marcher) with dt_gamma > 0 (step size grows with t): the differentiable
    forward still equals the inference renderer's, and the ray gradients still match the fp64 definition built from the stand-alone
    marcher's samples."""
    bitfield, grid = syn.stonehenge_bitfield(cascade=2)
    w = dict(all_active_weights())
    w2 = syn.ngp_weights(0, variant="B", bound=2)
    for k in ("offsets", "per_level_scale", "S", "bound"):
        w[k] = w2[k]
    rows = int(w["offsets"][-1])
    assert rows == w2["table"].shape[0]
    table = np.abs(w2["table_f32"]).astype(np.float16)
    w.update(table=table, table_f32=table.astype(np.float32))
    model = syn.build_model(w, bitfield, grid, device=cuda)
    assert model.cascade == 2
    o, d = syn.pinhole_rays(syn.orbit_pose(radius=2.6), 64, 64)
    pick = np.random.default_rng(4).permutation(o.shape[0])[:256]
    o = torch.from_numpy(np.ascontiguousarray(o[pick], np.float32)).to(cuda)
    d = torch.from_numpy(np.ascontiguousarray(d[pick], np.float32)).to(cuda)
    G = torch.randn(256, 3, generator=torch.Generator().manual_seed(6)).to(cuda)
    kw = dict(staged=True, bg_color=0.5, perturb=False, dt_gamma=1 / 128, max_steps=MAX_STEPS)
    with torch.no_grad():
        ref = model.render(o[None], d[None], **kw)
    orq, drq = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
    res = model.render(orq[None], drq[None], **kw)
    assert_same_render(res, ref)
    (res["image"][0] * G).sum().backward()

    N = o.shape[0]
    nears, fars = our_rm.near_far_from_aabb(o, d, model.aabb_infer, model.min_near)
    alive = torch.arange(N, dtype=torch.int32, device=cuda)
    _, _, deltas = our_rm.march_rays(N, MAX_STEPS, alive, nears.clone(), o, d, model.bound, model.density_bitfield, model.cascade,
                                     model.grid_size, nears, fars, -1, False, 1 / 128, MAX_STEPS)
    t, dt, count = dorc.samples_from_deltas(deltas, nears, N, MAX_STEPS)
    S = max(int(count.max()), 1)
    assert int((count > 0).sum()) > N // 4
    net = dorc.Net(w, device=cuda)
    od, dd = o.double().requires_grad_(True), d.double().requires_grad_(True)
    image, _ = dorc.render(net, od, dd, t[:, :S].contiguous(), dt[:, :S].contiguous(), count, bg=0.5)
    assert float((image.detach() - res["image"][0].double()).abs().max()) < 2e-3
    (image * G.double()).sum().backward()
    hit = count > 0
    mo, md = float(per_row_rel(orq.grad, od.grad)[hit].median()), float(per_row_rel(drq.grad, dd.grad)[hit].median())
    print(f"bound 2, dt_gamma 1/128: per-ray median error {mo:.3e} / {md:.3e}; aggregate {rel_l2(orq.grad, od.grad):.3e}")
    assert mo < 1e-2 and md < 1e-2
    assert rel_l2(orq.grad, od.grad) < 1e-1


@pytest.mark.parametrize("nl,nlc", [(3, 2), (4, 4)])
def test_other_network_depths(cuda, scene, nl, nlc):
    """num_layers / num_layers_color other than the default (2, 3): the rolled hidden-layer loops and the mask bookkeeping."""
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant="B", num_layers=nl, num_layers_color=nlc)
    model = syn.build_model(w, bitfield, grid, device=cuda)
    o, d = syn.pinhole_rays(syn.orbit_pose(), 64, 64)
    pick = np.random.default_rng(2).permutation(o.shape[0])[:128]
    o = torch.from_numpy(np.ascontiguousarray(o[pick], np.float32)).to(cuda)
    d = torch.from_numpy(np.ascontiguousarray(d[pick], np.float32)).to(cuda)
    G = torch.randn(128, 3, generator=torch.Generator().manual_seed(1)).to(cuda)
    with torch.no_grad():
        ref = model.render(o[None], d[None], staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS)
    res, go, gd = our_grads(model, o, d, G)
    assert_same_render(res, ref)
    t, dt, count = marched_samples(model, o, d)
    net = dorc.Net(w, device=cuda)
    od, dd = o.double().requires_grad_(True), d.double().requires_grad_(True)
    image, _ = dorc.render(net, od, dd, t, dt, count, bg=1.0)
    (image * G.double()).sum().backward()
    hit = count > 0
    mo, md = float(per_row_rel(go, od.grad)[hit].median()), float(per_row_rel(gd, dd.grad)[hit].median())
    print(f"layers {nl}/{nlc}: per-ray median error {mo:.3e} / {md:.3e}; aggregate {rel_l2(go, od.grad):.3e}")
    assert mo < 3e-2 and md < 3e-2 and rel_l2(go, od.grad) < 1.5e-1
    x = (torch.rand(2048, 3, generator=torch.Generator().manual_seed(5)) * 2 - 1).to(cuda)
    xr = x.clone().requires_grad_(True)
    model.density(xr)["sigma"].sum().backward()
    xd = x.double().requires_grad_(True)
    net.density(xd)[0].sum().backward()
    assert float(per_row_rel(xr.grad, xd.grad).median()) < 5e-3


def test_sample_budget(world, cuda):
    """max_steps = 6 (the step size is then dt_max = 2 sqrt(3) / 128, raymarching.cu:745-746): many rays run out of their sample budget
    inside the occupied region.  The warp-per-ray kernels give a ray exactly
    max_steps samples (the reference's loop, `while step < max_steps` over chunks of n_step, gives between max_steps and max_steps + 7
    depending on the batch), which is what the stand-alone marcher does with one chunk of n_step = max_steps."""
    model, w, o, d, G = world["mC"], world["wC"], world["o"], world["d"], world["G"]
    steps = 6
    orq, drq = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
    res = model.render(orq[None], drq[None], staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=steps)
    (res["image"][0] * G).sum().backward()
    N = o.shape[0]
    nears, fars = our_rm.near_far_from_aabb(o, d, model.aabb_infer, model.min_near)
    alive = torch.arange(N, dtype=torch.int32, device=cuda)
    _, _, deltas = our_rm.march_rays(N, steps, alive, nears.clone(), o, d, model.bound, model.density_bitfield, model.cascade,
                                     model.grid_size, nears, fars, -1, False, 0, steps)
    t, dt, count = dorc.samples_from_deltas(deltas, nears, N, steps)
    assert int((count == steps).sum()) > N // 8, "the budget should bind for many rays"
    net = dorc.Net(w, device=cuda)
    od, dd = o.double().requires_grad_(True), d.double().requires_grad_(True)
    image, ws = dorc.render(net, od, dd, t, dt, count, bg=1.0)
    assert float((image.detach() - res["image"][0].double()).abs().max()) < 2e-3
    assert float((ws.detach() - res["weights_sum"].double()).abs().max()) < 2e-3
    (image * G.double()).sum().backward()
    hit = count > 0
    mo = float(per_row_rel(orq.grad, od.grad)[hit].median())
    print(f"max_steps {steps}: per-ray median error {mo:.3e}, aggregate {rel_l2(orq.grad, od.grad):.3e}")
    assert mo < 5e-3 and rel_l2(orq.grad, od.grad) < 5e-2


def reference_get_rays(poses, intrinsics, H, W):
    """nerf/utils.py:52-116 with N = -1, restated in torch (differentiable w.r.t. the poses, as the reference's is)."""
    fx, fy, cx, cy = intrinsics
    device = poses.device
    i, j = torch.meshgrid(torch.linspace(0, W - 1, W, device=device), torch.linspace(0, H - 1, H, device=device), indexing="ij")
    i = i.t().reshape(1, H * W) + 0.5
    j = j.t().reshape(1, H * W) + 0.5
    zs = torch.ones_like(i)
    directions = torch.stack(((i - cx) / fx * zs, (j - cy) / fy * zs, zs), dim=-1)
    directions = directions / torch.norm(directions, dim=-1, keepdim=True)
    rays_d = directions @ poses[:, :3, :3].transpose(-1, -2)
    rays_o = poses[..., :3, 3][..., None, :].expand_as(rays_d)
    return rays_o, rays_d


def test_pose_gradient_through_get_rays_and_render(world, cuda):
    """Estimator.measurement_fn's chain (nav/estimator_helpers.py:203-218): pose -> get_rays -> pixel batch -> render_fn -> loss.backward().
    get_rays carries the pose gradient like the reference's torch version; the whole chain's d loss / d pose equals the one obtained
    with the torch restatement of get_rays in front of the same renderer."""
    from nerfsafetyvalidation_b200.nerf.utils import get_rays
    model = world["mB"]
    H = W = 40
    intr = (55.0, 55.0, W / 2, H / 2)
    pose0 = torch.from_numpy(syn.orbit_pose()[None].astype(np.float32)).to(cuda)
    batch = torch.from_numpy(np.random.default_rng(1).permutation(H * W)[:512]).to(cuda)
    target = torch.rand(512, 3, generator=torch.Generator().manual_seed(2)).to(cuda)
    grads, images = [], []
    for rays_fn in (lambda p: tuple(get_rays(p, intr, H, W)[k] for k in ("rays_o", "rays_d")), lambda p: reference_get_rays(p, intr, H, W)):
        pose = pose0.clone().requires_grad_(True)
        ro, rd = rays_fn(pose)
        assert ro.requires_grad and rd.requires_grad
        out = model.render(ro[:, batch], rd[:, batch], staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS)
        loss = torch.nn.functional.mse_loss(out["image"].reshape(-1, 3), target)
        loss.backward()
        grads.append(pose.grad.clone())
        images.append(out["image"].detach())
    assert float(grads[0][:, :3].abs().max()) > 0 and float(grads[0][:, 3].abs().max()) == 0
    assert float((images[0] - images[1]).abs().max()) < 1e-4
    # the rays of the two versions agree to ~2e-7 (test_gpu_parity.py), so the renderer sees nearly the same inputs
    assert rel_l2(grads[0], grads[1]) < 2e-2, (grads[0], grads[1])


def blob_scene_weights():
    """A hand-built NGP network that behaves like a TRAINED scene where it matters for pose refinement: a smooth density blob that has
    faded to ~1 % of its peak where the occupancy grid cuts it off, smooth colour, nothing at the scale of the fine hash levels.
    Level 0 (dense 17^3 grid) carries f0 = exp(-|p|^2 / (2 * 0.35^2)) and f1 = 0.5 + 0.5 sin(3x) cos(2y); level 1 is the constant 1 (the
    networks have no biases); every other level is zero.  sigma = exp(12 f0 - 8), rgb = sigmoid((4 f1 - 2, 2 - 4 f1, 0.5))."""
    w = dict(syn.ngp_weights(0, variant="A"))
    table = np.zeros_like(w["table_f32"])
    off = np.asarray(w["offsets"], np.int64)
    i = np.arange(17, dtype=np.float64)
    p = ((i - 0.5) / 15.0) * 2 - 1                                  # world coordinate of level-0 grid node i (pos = u * 15 + 0.5)
    Z, Y, X = np.meshgrid(p, p, p, indexing="ij")                   # row = x + 17 y + 289 z
    f0 = np.exp(-(X * X + Y * Y + Z * Z) / (2 * 0.35 ** 2))
    f1 = 0.5 + 0.5 * np.sin(3 * X) * np.cos(2 * Y)
    table[off[0]:off[0] + 17 ** 3, 0] = f0.reshape(-1)
    table[off[0]:off[0] + 17 ** 3, 1] = f1.reshape(-1)
    table[off[1]:off[2], 0] = 1.0
    ws = np.zeros(64 * (32 + 64 + 16), np.float32)
    W0, W1, W2 = ws[:2048].reshape(64, 32), ws[2048:2048 + 4096].reshape(64, 64), ws[2048 + 4096:].reshape(16, 64)
    W0[0, 0] = W0[1, 1] = W0[2, 2] = 1.0                            # hidden units 0, 1, 2 <- f0, f1, 1
    W1[0, 0] = W1[1, 1] = W1[2, 2] = 1.0
    W2[0, 0], W2[0, 2] = 12.0, -8.0                                 # log-density
    W2[1, 1] = 1.0                                                  # geo_feat[0] = f1
    W2[2, 2] = 1.0                                                  # geo_feat[1] = 1
    wc = np.zeros(64 * (32 + 2 * 64 + 16), np.float32)
    C0, C1, C2, C3 = (wc[:2048].reshape(64, 32), wc[2048:2048 + 4096].reshape(64, 64), wc[2048 + 4096:2048 + 8192].reshape(64, 64),
                      wc[2048 + 8192:].reshape(16, 64))
    C0[0, 16] = C0[1, 17] = 1.0                                     # colour input columns 16, 17 = geo_feat[0], geo_feat[1]
    C1[0, 0] = C1[1, 1] = C2[0, 0] = C2[1, 1] = 1.0
    C3[0, 0], C3[0, 1] = 4.0, -2.0
    C3[1, 0], C3[1, 1] = -4.0, 2.0
    C3[2, 1] = 0.5
    t16, ws16, wc16 = table.astype(np.float16), ws.astype(np.float16), wc.astype(np.float16)
    w.update(table=t16, table_f32=t16.astype(np.float32), w_sigma=ws16, w_sigma_f32=ws16.astype(np.float32), w_color=wc16,
             w_color_f32=wc16.astype(np.float32), density_scale=1.0)
    return w


def test_pose_refinement_on_a_smooth_scene(cuda):
    """The estimator's loop in miniature (nav/estimator_helpers.py:120-160): the observation is the render at the true pose, the state
    starts 3 cm off, Adam follows d loss / d pose through get_rays and the differentiable render.  The occupancy grid is derived from
    the network itself by update_extra_state (renderer.py:451-546).  The loss and the pose error both shrink."""
    from nerfsafetyvalidation_b200.nerf.utils import get_rays
    w = blob_scene_weights()
    model = syn.build_model(w, None, None, device=cuda)
    torch.manual_seed(0)
    model.update_extra_state()
    occupied = int((model.density_grid > min(model.mean_density, model.density_thresh)).sum())
    assert 1000 < occupied < 128 ** 3 // 4, occupied
    H = W = 48
    intr = (60.0, 60.0, W / 2, H / 2)
    true_pose = torch.from_numpy(syn.orbit_pose()[None].astype(np.float32)).to(cuda)
    kw = dict(staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS)
    with torch.no_grad():
        r = get_rays(true_pose, intr, H, W)
        target = model.render(r["rays_o"], r["rays_d"], **kw)["image"][0]
    assert float(target.std()) > 0.02, "the blob should be visible"
    batch = torch.from_numpy(np.random.default_rng(3).permutation(H * W)[:1024]).to(cuda)
    delta0 = torch.tensor([0.02, -0.015, 0.015], device=cuda)
    delta = delta0.clone().requires_grad_(True)
    opt = torch.optim.Adam([delta], lr=1e-3)
    losses = []
    with model.frozen_weights():
        for _ in range(80):
            pose = true_pose.clone()
            pose[:, :3, 3] = true_pose[:, :3, 3] + delta
            rays = get_rays(pose, intr, H, W)
            out = model.render(rays["rays_o"][:, batch], rays["rays_d"][:, batch], **kw)
            loss = torch.nn.functional.mse_loss(out["image"][0], target[batch])
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(float(loss))
    err0, err1 = float(delta0.norm()), float(delta.detach().norm())
    print(f"pose refinement: loss {losses[0]:.3e} -> {losses[-1]:.3e}, translation error {err0:.4f} -> {err1:.4f}")
    assert losses[-1] < 0.25 * losses[0]
    assert err1 < 0.5 * err0


def test_density_descent_on_a_smooth_scene(cuda):
    """The planner's use of density_fn under autograd (nav/quad_plot.py:223-300: way-points are pushed out of dense regions by gradient
    descent on a collision cost): on the blob scene, descending sum(sigma) moves every point away from the blob's centre."""
    model = syn.build_model(blob_scene_weights(), None, None, device=cuda)
    g = torch.Generator().manual_seed(4)
    dirs = torch.randn(256, 3, generator=g)
    x0 = (dirs / dirs.norm(dim=-1, keepdim=True) * (0.15 + 0.2 * torch.rand(256, 1, generator=g))).to(cuda)
    x = x0.clone().requires_grad_(True)
    opt = torch.optim.SGD([x], lr=1e-3)
    with model.frozen_weights():
        first = None
        for _ in range(30):
            cost = model.density(x)["sigma"].sum()
            first = float(cost) if first is None else first
            opt.zero_grad()
            cost.backward()
            opt.step()
    last = float(model.density(x.detach())["sigma"].sum())
    moved_out = (x.detach().norm(dim=-1) > x0.norm(dim=-1)).float().mean()
    print(f"density descent: cost {first:.3e} -> {last:.3e}, {float(moved_out) * 100:.0f} % of the points moved outwards")
    assert last < 0.7 * first and float(moved_out) > 0.95


def test_hessian_call_of_the_estimator(cuda):
    """nav/estimator_helpers.py:297 takes torch.autograd.functional.hessian of measurement_fn.  Upstream the render's backward kernels
    return plain tensors (their @once_differentiable is commented out), so the second differentiation sees the photometric term's
    gradient as a constant and the Hessian is that of the torch-side process term alone; the same call must run here, with the same
    outcome."""
    from nerfsafetyvalidation_b200.nerf.utils import get_rays
    model = syn.build_model(blob_scene_weights(), None, None, device=cuda)
    torch.manual_seed(0)
    model.update_extra_state()
    H = W = 24
    intr = (30.0, 30.0, W / 2, H / 2)
    true_pose = torch.from_numpy(syn.orbit_pose()[None].astype(np.float32)).to(cuda)
    target = torch.rand(H * W, 3, generator=torch.Generator().manual_seed(1)).to(cuda)

    def measurement(delta):
        pose = true_pose.clone()
        pose[:, :3, 3] = true_pose[:, :3, 3] + delta
        rays = get_rays(pose, intr, H, W)
        out = model.render(rays["rays_o"], rays["rays_d"], staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=MAX_STEPS)
        return torch.nn.functional.mse_loss(out["image"][0], target) + 3.0 * (delta * delta).sum()

    with model.frozen_weights():
        hess = torch.autograd.functional.hessian(measurement, torch.zeros(3, device=cuda))
        grad = torch.autograd.functional.jacobian(measurement, torch.zeros(3, device=cuda))
    assert torch.allclose(hess, 6.0 * torch.eye(3, device=cuda), atol=1e-6), hess
    assert float(grad.abs().max()) > 0          # the first derivative does see the photometric term
