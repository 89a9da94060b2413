"""GPU parity tests (run on the B200 box: `pytest -m gpu`).  Every call goes through the C ABI (ctypes) of
libngp_b200.so.  Oracles, strongest first:
  * the reference's own extensions rebuilt for sm_100a (tests/refext.py -> oracle/_ref/*.so): bit-exact for march /
    composite / near-far / packbits / morton / grid (reference rounding), tolerance for the fp32-accumulate MLP;
  * the C restatement (oracle/ngp_oracle.c): integer outputs exact, floats as documented per test.
SURVEY.md Appendix C is the matrix these follow.
"""
import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import _cabi, synthetic as syn
from nerfsafetyvalidation_b200 import ffmlp as our_ffmlp
from nerfsafetyvalidation_b200 import gridencoder as our_grid
from nerfsafetyvalidation_b200 import raymarching as our_rm
from nerfsafetyvalidation_b200 import shencoder as our_sh
from oracle import oracle as orc
from tests import refext
from tests.inputs import grid_inputs, make_rays

pytestmark = pytest.mark.gpu

needs_ref = pytest.mark.skipif(not refext.available(), reason="oracle/_ref/*.so not built (python oracle/build_ref.py)")


def bits(t):
    return t.contiguous().view(torch.int32)


def assert_bitwise(a, b, what):
    a, b = a.contiguous(), b.contiguous()
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    neq = bits(a) != bits(b)
    n = int(neq.sum().item())
    assert n == 0, f"{what}: {n} of {neq.numel()} words differ; first at {neq.flatten().nonzero()[:4].flatten().tolist()}"


@pytest.fixture(scope="module")
def frame(cuda, scene):
    """cfg2-shaped camera at reduced resolution + models (ours, reference) for both weight variants."""
    bitfield, grid = scene
    pose = syn.orbit_pose()
    o, d = syn.pinhole_rays(pose, 160, 160)
    out = dict(o=torch.from_numpy(o).to(cuda), d=torch.from_numpy(d).to(cuda), bitfield=bitfield, grid=grid)
    for var in "AB":
        w = syn.ngp_weights(0, variant=var)
        out["w" + var] = w
        out["ours" + var] = syn.build_model(w, bitfield, grid, device=cuda)
        if refext.available():
            out["ref" + var] = refext.RefModel(w, bitfield, device=cuda)
    return out


# ---------------------------------------------------------------------------------------------- setup kernels
def test_library_loaded_and_counts_launches(cuda):
    lib = _cabi.load()
    assert lib.ngp_abi_version() == 1
    before = _cabi.launch_count()
    our_rm.morton3D(torch.zeros(4, 3, dtype=torch.int32, device=cuda))
    assert _cabi.launch_count() == before + 1


def test_cpu_tensor_is_an_error_not_a_fallback(cuda):
    lib = _cabi.load()
    host = torch.zeros(8, 3)
    dev = torch.zeros(8, dtype=torch.int32, device=cuda)
    rc = lib.ngp_morton3D(_cabi.ptr(host.int()), 8, _cabi.ptr(dev), None)
    assert rc != 0 and "device pointer" in _cabi.last_error()


def test_morton_all_cells(cuda):
    idx = torch.arange(128, dtype=torch.int32)
    coords = torch.stack(torch.meshgrid(idx, idx, idx, indexing="ij"), -1).reshape(-1, 3).to(cuda)
    ours = our_rm.morton3D(coords)
    want = orc.morton3D(coords.cpu().numpy())
    assert np.array_equal(ours.cpu().numpy(), want)
    assert sorted(ours.cpu().tolist()) == list(range(128 ** 3))
    back = our_rm.morton3D_invert(ours)
    assert torch.equal(back, coords)
    if refext.available():
        assert torch.equal(ours, refext.morton3D(coords))
        assert torch.equal(back, refext.morton3D_invert(ours))
    rnd = torch.randint(0, 2 ** 30, (100000,), dtype=torch.int32, device=cuda)
    assert np.array_equal(our_rm.morton3D_invert(rnd).cpu().numpy(), orc.morton3D_invert(rnd.cpu().numpy()))


def test_packbits(cuda, scene):
    bitfield, grid = scene
    g = torch.from_numpy(grid).to(cuda)
    thresh = min(float(np.mean(np.clip(grid, 0, None))), 10.0)
    ours = our_rm.packbits(g, thresh)
    assert np.array_equal(ours.cpu().numpy(), bitfield)
    assert np.array_equal(ours.cpu().numpy(), orc.packbits(grid, thresh))
    gen = torch.Generator().manual_seed(1)
    rnd = torch.rand(2, 128 ** 3, generator=gen).to(cuda)
    rnd[0, ::7] = 0.5       # values exactly at the threshold are NOT set (strict >)
    for t in (0.5, -1.0, 2.0):
        ours = our_rm.packbits(rnd, t)
        assert np.array_equal(ours.cpu().numpy(), orc.packbits(rnd.cpu().numpy(), t))
        if refext.available():
            assert torch.equal(ours, refext.packbits(rnd, t))


def test_near_far_bitwise(cuda):
    o, d = make_rays(65536, 0, cuda)
    for b, min_near in ((1.0, 0.2), (2.0, 0.05)):
        aabb = torch.tensor([-b, -b, -b, b, b, b], dtype=torch.float32, device=cuda)
        n, f = our_rm.near_far_from_aabb(o, d, aabb, min_near)
        cn, cf = orc.near_far_from_aabb(o.cpu().numpy(), d.cpu().numpy(), aabb.cpu().numpy(), min_near)
        assert_bitwise(n, torch.from_numpy(cn).to(cuda), "nears vs C oracle")
        assert_bitwise(f, torch.from_numpy(cf).to(cuda), "fars vs C oracle")
        if refext.available():
            rn, rf = refext.near_far_from_aabb(o, d, aabb, min_near)
            assert_bitwise(n, rn, "nears vs reference ext")
            assert_bitwise(f, rf, "fars vs reference ext")
        miss = bits(n) == 0x7F7FFFFF                      # FLT_MAX bit pattern
        assert torch.equal(miss, bits(f) == 0x7F7FFFFF)
        if b == 1.0:   # with b = 2 every origin lies inside the box
            assert 0 < int(miss.sum()) < n.numel()


def test_compact_alive(cuda):
    gen = torch.Generator().manual_seed(3)
    for n in (1, 2, 31, 32, 33, 2047, 2048, 2049, 100003, 2_000_000):
        x = torch.arange(n, dtype=torch.int32)
        x[torch.rand(n, generator=gen) < 0.37] = -1
        xd = x.to(cuda)
        out = our_rm.compact_alive(xd)
        assert torch.equal(out, xd[xd >= 0]), f"n={n}"
    none = torch.full((5000,), -1, dtype=torch.int32, device=cuda)
    assert our_rm.compact_alive(none).numel() == 0
    assert our_rm.compact_alive(torch.empty(0, dtype=torch.int32, device=cuda)).numel() == 0


# ---------------------------------------------------------------------------------------------- march / composite
@needs_ref
@pytest.mark.parametrize("variant", ["A", "B"])
def test_march_and_composite_teacher_forced(frame, cuda, variant):
    """At snapshots of the reference loop: same (rays_alive, rays_t) in -> bitwise xyzs/dirs/deltas out; then the
    reference's own sigmas/rgbs/deltas in -> exact alive flags, bitwise rays_t and accumulators out."""
    ref = frame["ref" + variant]
    rec = []
    ref.run_cuda(frame["o"], frame["d"], record=rec)
    assert len(rec) > 10
    picks = sorted({0, 1, 5, min(50, len(rec) - 1), len(rec) // 2, len(rec) - 1})
    nears, fars = our_rm.near_far_from_aabb(frame["o"], frame["d"], ref.aabb, ref.min_near)
    seen_steps = set()
    for i in picks:
        r = rec[i]
        seen_steps.add(r["n_step"])
        xyzs, dirs, deltas = our_rm.march_rays(r["n_alive"], r["n_step"], r["rays_alive"], r["rays_t"], frame["o"], frame["d"],
                                                ref.bound, ref.bitfield, 1, 128, nears, fars, 128, False, 0, 1024)
        assert_bitwise(xyzs, r["xyzs"], f"iter {i} xyzs")
        assert_bitwise(dirs, r["dirs"], f"iter {i} dirs")
        assert_bitwise(deltas, r["deltas"], f"iter {i} deltas")
        alive = r["rays_alive"].clone()
        rays_t, ws, dp, im = r["rays_t"].clone(), r["weights_sum"].clone(), r["depth"].clone(), r["image"].clone()
        our_rm.composite_rays(r["n_alive"], r["n_step"], alive, rays_t, r["sigmas"], r["rgbs"], r["deltas"], ws, dp, im)
        assert torch.equal(alive, r["alive_after"]), f"iter {i} alive flags"
        assert_bitwise(rays_t, r["rays_t_after"], f"iter {i} rays_t")
        assert_bitwise(ws, r["weights_sum_after"], f"iter {i} weights_sum")
        assert_bitwise(dp, r["depth_after"], f"iter {i} depth")
        assert_bitwise(im, r["image_after"], f"iter {i} image")
        assert torch.equal(our_rm.compact_alive(alive), alive[alive >= 0])
    assert len(seen_steps) >= 2


@needs_ref
@pytest.mark.parametrize("bound,dt_gamma,n_step", [(1.0, 0.0, 1), (1.0, 0.0, 2), (1.0, 0.0, 8), (1.0, 1.0 / 128, 4), (2.0, 0.0, 8),
                                                   (2.0, 1.0 / 128, 3), (4.0, 1.0 / 256, 8)])
def test_march_bitwise_random_rays(cuda, scene, bound, dt_gamma, n_step):
    """random rays, cascades 1..3, exponential stepping on/off, dense / sparse / empty bitfields"""
    cascade = 1 + int(np.ceil(np.log2(bound)))
    o, d = make_rays(20000, 7, cuda)
    o = o * bound / 2
    aabb = torch.tensor([-bound] * 3 + [bound] * 3, dtype=torch.float32, device=cuda)
    nears, fars = our_rm.near_far_from_aabb(o, d, aabb, 0.05)
    gen = torch.Generator().manual_seed(11)
    sparse = torch.from_numpy(np.tile(scene[0], cascade)).to(cuda)
    fields = {"sparse": sparse, "dense": torch.full_like(sparse, 255), "empty": torch.zeros_like(sparse),
              "random": torch.randint(0, 256, sparse.shape, generator=gen, dtype=torch.uint8).to(cuda)}
    alive = torch.randperm(20000, generator=gen)[:15000].int().to(cuda)
    rays_t = nears + torch.rand(20000, generator=gen).to(cuda) * 0.3
    for name, bf in fields.items():
        a = our_rm.march_rays(15000, n_step, alive, rays_t, o, d, bound, bf, cascade, 128, nears, fars, 128, False, dt_gamma, 1024)
        b = refext.march_rays(15000, n_step, alive, rays_t, o, d, bound, bf, cascade, 128, nears, fars, 128, False, dt_gamma, 1024)
        for x, y, what in zip(a, b, ("xyzs", "dirs", "deltas")):
            assert_bitwise(x, y, f"{name} {what}")
        co = orc.march_rays(15000, n_step, alive.cpu().numpy(), rays_t.cpu().numpy(), o.cpu().numpy(), d.cpu().numpy(), bound,
                            bf.cpu().numpy(), cascade, 128, nears.cpu().numpy(), fars.cpu().numpy(), 128, dt_gamma, 1024)
        for x, y, what in zip(a, co, ("xyzs", "dirs", "deltas")):
            assert_bitwise(x, torch.from_numpy(y).to(cuda), f"{name} {what} vs C oracle")
    assert int((a[2][:, 0] != 0).sum()) > 0


# ---------------------------------------------------------------------------------------------- encoders / MLP
@pytest.mark.parametrize("log2T", [19, 14])
def test_grid_indices_exact_and_features(cuda, log2T):
    w = syn.ngp_weights(0, log2_hashmap_size=log2T, variant="B")
    x = grid_inputs(50000, 5, cuda)
    lib = _cabi.load()
    offsets = torch.from_numpy(w["offsets"]).to(cuda)
    B, L = x.shape[0], 16
    scales = torch.empty(L, dtype=torch.float32, device=cuda)
    _cabi.check(lib.ngp_grid_level_scales(L, w["S"], 16, _cabi.ptr(scales), _cabi.stream()))
    scales_h = scales.cpu().numpy()
    for dtype, key in ((torch.float16, "table"), (torch.float32, "table_f32")):
        table = torch.from_numpy(w[key]).to(cuda)
        for ref_rounding in (1, 0):
            out = torch.empty(B, L * 2, dtype=dtype, device=cuda)
            idx = torch.empty(L, B, 8, dtype=torch.int32, device=cuda)
            _cabi.check(lib.ngp_grid_encode_forward(_cabi.ptr(x), _cabi.ptr(table), 1 if dtype == torch.float16 else 0, _cabi.ptr(offsets),
                                                    _cabi.ptr(out), B, 3, 2, L, w["S"], 16, 0, 0, 1, ref_rounding, _cabi.ptr(idx),
                                                    _cabi.stream()))
            c_out, c_idx = orc.grid_encode(x.cpu().numpy(), w[key], w["offsets"], w["S"], 16, scales=scales_h, round_mode=ref_rounding,
                                           want_indices=True)
            assert np.array_equal(idx.cpu().numpy(), c_idx), "gathered table rows differ from the C oracle"
            got = out.cpu().numpy()
            if dtype == torch.float16:
                assert np.array_equal(got.view(np.uint16), c_out.view(np.uint16)), f"half features (round_mode {ref_rounding}) vs C oracle"
            else:
                assert np.array_equal(got.view(np.uint32), c_out.view(np.uint32)), "fp32 features vs C oracle"
            if refext.available() and (ref_rounding == 1 or dtype == torch.float32):
                r = refext.grid_encode(x, table, offsets, w["per_level_scale"])
                assert_bitwise(out.float(), r.float(), f"features vs reference ext ({dtype})")
    assert (got[128:256] == 0).all() and np.abs(got[256:]).max() > 0


def test_grid_encoder_module_matches_c_oracle(cuda):
    enc = our_grid.GridEncoder(desired_resolution=2048).to(cuda)
    with torch.no_grad():
        enc.embeddings.uniform_(-1, 1)
    x = (grid_inputs(4096, 9, cuda) * 2 - 1)
    with torch.no_grad():
        y32 = enc(x, bound=1)
        with torch.autocast("cuda", dtype=torch.float16):
            y16 = enc(x, bound=1)
    assert y32.dtype == torch.float32 and y16.dtype == torch.float16 and y32.shape == (4096, 32)
    x01 = ((x + 1) / 2).cpu().numpy()
    lib = _cabi.load()
    scales = torch.empty(16, dtype=torch.float32, device=cuda)
    S = float(np.log2(enc.per_level_scale))
    _cabi.check(lib.ngp_grid_level_scales(16, S, 16, _cabi.ptr(scales), _cabi.stream()))
    c32, _ = orc.grid_encode(x01, enc.embeddings.detach().cpu().numpy(), enc.offsets.cpu().numpy(), S, 16, scales=scales.cpu().numpy())
    assert np.array_equal(y32.cpu().numpy().view(np.uint32), c32.view(np.uint32))
    c16, _ = orc.grid_encode(x01, enc.embeddings.detach().half().cpu().numpy(), enc.offsets.cpu().numpy(), S, 16, scales=scales.cpu().numpy(),
                             round_mode=1)
    assert np.array_equal(y16.cpu().numpy().view(np.uint16), c16.view(np.uint16))


def test_sh_encode(cuda):
    g = torch.Generator().manual_seed(2)
    d = torch.randn(30000, 3, generator=g)
    d[:20000] = d[:20000] / d[:20000].norm(dim=-1, keepdim=True)
    d = d.to(cuda)
    for degree in (1, 2, 3, 4):
        ours = our_sh.sh_encode(d, degree)
        want = orc.sh_encode(d.cpu().numpy(), degree)
        assert np.allclose(ours.cpu().numpy(), want, atol=1e-6 * max(1.0, float(np.abs(want).max())), rtol=1e-6)
    if refext.available():
        ref = refext.sh_encode(d, 4)
        assert torch.allclose(our_sh.sh_encode(d, 4), ref, atol=1e-5, rtol=1e-6)
    mod = our_sh.SHEncoder()
    assert mod(d.view(300, 100, 3)).shape == (300, 100, 16)


def torch_mlp_chain(x_h, w_h, in_dim, num_layers, out_dim=16):
    """fp32 `relu(x @ W^T)` chain with half storage between layers (what fp32-accumulate tensor cores compute)."""
    x = x_h.float()
    off = 0
    dims = [in_dim] + [64] * num_layers
    for i in range(num_layers):
        W = w_h[off: off + 64 * dims[i]].float().view(64, dims[i])
        off += 64 * dims[i]
        x = torch.relu(x @ W.t()).half().float()
    W = w_h[off: off + out_dim * 64].float().view(out_dim, 64)
    return (x @ W.t()).half()


@pytest.mark.parametrize("in_dim,num_layers,B", [(32, 2, 128), (32, 3, 1000), (32, 2, 65536 + 17), (64, 4, 4096), (16, 2, 300)])
def test_ffmlp(cuda, in_dim, num_layers, B):
    g = torch.Generator().manual_seed(in_dim + num_layers)
    x = (torch.randn(B, in_dim, generator=g) * 0.5).half().to(cuda)
    mlp = our_ffmlp.FFMLP(in_dim, 16, 64, num_layers).to(cuda).eval()
    with torch.no_grad():
        y = mlp(x)
    assert y.shape == (B, 16) and y.dtype == torch.float16
    wh = mlp.weights.detach().half()
    want = torch_mlp_chain(x, wh, in_dim, num_layers)
    err = (y.float() - want.float()).abs().max().item()
    assert err <= 1e-3 * max(1.0, want.float().abs().max().item()), f"vs fp32 torch chain: {err}"
    c = orc.ffmlp(x.cpu().numpy(), wh.cpu().numpy(), in_dim, 16, 64, num_layers)
    errc = np.abs(y.cpu().numpy().astype(np.float32) - c.astype(np.float32)).max()
    assert errc <= 1e-3 * max(1.0, float(np.abs(c.astype(np.float32)).max())), f"vs C oracle: {errc}"
    if refext.available():
        r = refext.ffmlp(x, wh, in_dim, 16, 64, num_layers)
        errr = (y.float() - r.float()).abs().max().item()
        assert errr <= 4e-3 * max(1.0, r.float().abs().max().item()), f"vs reference ext (fp16 accumulate): {errr}"


@pytest.mark.parametrize("variant", ["A", "B"])
def test_network_forward(frame, cuda, variant):
    w, ours = frame["w" + variant], frame["ours" + variant]
    g = torch.Generator().manual_seed(4)
    x = (torch.rand(20000, 3, generator=g) * 2 - 1).to(cuda)
    x[:16] = 0.0          # padded rows of the march output are all-zero and still evaluated (raymarching.py:326-331)
    d = torch.randn(20000, 3, generator=g)
    d = (d / d.norm(dim=-1, keepdim=True)).to(cuda)
    d[:16] = 0.0
    with torch.no_grad():
        sig, rgb = ours(x, d)
        sig_u, rgb_u = ours._forward_unfused(x, d)
    assert sig.shape == (20000,) and rgb.shape == (20000, 3)
    # the default engine is tcgen05 (TMEM accumulators); the mma.sync formulation of the same MLPs must agree bit for bit: both
    # accumulate each layer in fp32 over K in the same 16-wide steps and round to fp16 at the same points
    ours.mlp_engine = "mma_sync"
    try:
        with torch.no_grad():
            sig_m, rgb_m = ours(x, d)
    finally:
        ours.mlp_engine = "tcgen05"
    assert ((sig - sig_m).abs() / sig_m.abs().clamp_min(1e-6)).max().item() <= 1e-6 and (rgb - rgb_m).abs().max().item() <= 1e-6
    for M in (1, 127, 129):   # ragged tiles
        with torch.no_grad():
            s1, c1 = ours(x[:M], d[:M])
        assert torch.equal(s1, sig[:M]) and torch.equal(c1, rgb[:M])
    lib = _cabi.load()
    scales = torch.empty(16, dtype=torch.float32, device=cuda)
    _cabi.check(lib.ngp_grid_level_scales(16, w["S"], 16, _cabi.ptr(scales), _cabi.stream()))
    csig, crgb = orc.network_forward(x.cpu().numpy(), d.cpu().numpy(), w["table"], w["offsets"], w["S"], 16, w["w_sigma"], w["w_color"],
                                     scales=scales.cpu().numpy())
    rel = np.abs(sig.cpu().numpy() - csig) / np.maximum(np.abs(csig), 1e-6)
    assert rel.max() <= 1e-2, f"sigma vs C oracle rel {rel.max()}"
    assert np.abs(rgb.cpu().numpy() - crgb).max() <= 2e-3
    # unfused composition of our stand-alone ops (fp32 table outside autocast) agrees to tolerance
    assert ((sig - sig_u).abs() / sig_u.abs().clamp_min(1e-6)).max().item() <= 2e-2
    assert (rgb - rgb_u.float()).abs().max().item() <= 4e-3
    if refext.available():
        rs, rr = frame["ref" + variant].forward(x, d)
        assert ((sig - rs).abs() / rs.abs().clamp_min(1e-6)).max().item() <= 2e-2, "sigma vs reference ext"
        assert (rgb - rr.float()).abs().max().item() <= 4e-3, "rgb vs reference ext"


@pytest.mark.parametrize("variant", ["A", "B"])
def test_density_query(frame, cuda, variant):
    """NeRFNetwork.density (network_ff.py:76-89): the fused kernel against forward()'s sigma (identical
    arithmetic: bit-exact), against the composition of the stand-alone ops, and against the reference extensions."""
    ours = frame["ours" + variant]
    g = torch.Generator().manual_seed(5)
    x = (torch.rand(30001, 3, generator=g) * 2 - 1).to(cuda)
    x[:8] = 0.0
    x[8:16] = 1.5         # outside the box: the encoding is all zeros (gridencoder.cu:99-123)
    d = torch.zeros_like(x)
    d[:, 2] = 1.0
    before = _cabi.launch_count()
    with torch.no_grad():
        out = ours.density(x)
        sig_f, _ = ours(x, d)
    assert _cabi.launch_count() - before == 2                      # one kernel per call
    assert out["sigma"].shape == (30001,) and out["geo_feat"].shape == (30001, 15) and out["geo_feat"].dtype == torch.float16
    assert torch.equal(out["sigma"], sig_f)
    ours.fused_density = False
    try:
        with torch.no_grad():
            u = ours.density(x)
    finally:
        ours.fused_density = True
    assert ((out["sigma"] - u["sigma"].float()).abs() / u["sigma"].float().abs().clamp_min(1e-6)).max().item() <= 2e-2
    gmax = max(1.0, u["geo_feat"].float().abs().max().item())
    assert (out["geo_feat"].float() - u["geo_feat"].float()).abs().max().item() <= 4e-3 * gmax
    for M in (0, 1, 127, 129):
        with torch.no_grad():
            o = ours.density(x[:M])
        assert torch.equal(o["sigma"], out["sigma"][:M]) and torch.equal(o["geo_feat"], out["geo_feat"][:M])
    with torch.no_grad():
        o3 = ours.density(x[:600].view(20, 30, 3))                # leading dimensions are kept, like the reference's
    assert o3["sigma"].shape == (20, 30) and o3["geo_feat"].shape == (20, 30, 15)
    if refext.available():
        rs, _ = frame["ref" + variant].forward(x, d)
        assert ((out["sigma"] - rs).abs() / rs.abs().clamp_min(1e-6)).max().item() <= 2e-2, "sigma vs reference ext"


# ---------------------------------------------------------------------------------------------- end to end
def render_ours(model, o, d, **kw):
    with torch.no_grad():
        return model.render(o[None], d[None], bg_color=1, perturb=False, return_stats=True, **kw)


@needs_ref
@pytest.mark.parametrize("variant", ["A", "B"])
def test_render_vs_reference_loop(frame, cuda, variant):
    ours, ref = frame["ours" + variant], frame["ref" + variant]
    o, d = frame["o"], frame["d"]
    r = ref.run_cuda(o, d)
    out = render_ours(ours, o, d)
    img, dep = out["image"][0], out["depth"][0]
    assert img.shape == (o.shape[0], 3) and dep.shape == (o.shape[0],)
    miss = bits(r["nears"]) == 0x7F7FFFFF
    assert torch.equal(torch.isnan(dep), torch.isnan(r["depth"])), "NaN depth pattern (missed rays: 0/0, renderer.py:376)"
    assert torch.equal(torch.isnan(dep), miss)
    assert (img - r["image"]).abs().max().item() <= 1e-3
    assert (dep - r["depth"])[~miss].abs().max().item() <= 1e-3
    assert (out["weights_sum"] - r["weights_sum"]).abs().max().item() <= 1e-3
    stats = out["stats"].cpu().tolist()
    spr_diff = (out["samples_per_ray"] != r["samples_per_ray"])
    n_diff = int(spr_diff.sum())
    if variant == "A":   # nothing terminates by transmittance: the integer outputs must be identical
        assert n_diff == 0 and stats[0] == r["iterations"] and stats[1] == r["samples"]
    else:                # rays sitting on the T < 1e-4 threshold may flip with fp32- vs fp16-accumulated sigma
        assert n_diff <= max(4, o.shape[0] // 500), f"{n_diff} rays differ in sample count"
        assert abs(stats[0] - r["iterations"]) <= 2
    # stepwise mode (reference loop on our stand-alone ops) agrees with the fused launch
    with torch.no_grad():
        sw = ours.render(o[None], d[None], bg_color=1, perturb=False, fused=False)
    assert (sw["image"][0] - img).abs().max().item() <= 1e-3


@pytest.mark.parametrize("variant", ["A", "B"])
def test_render_vs_c_oracle(frame, cuda, variant):
    """Small tile so the CPU loop finishes in seconds; same fp32-accumulate contract as the fused kernel."""
    w, ours = frame["w" + variant], frame["ours" + variant]
    o, d = syn.pinhole_rays(syn.orbit_pose(110.0), 48, 48)
    lib = _cabi.load()
    scales = torch.empty(16, dtype=torch.float32, device=cuda)
    _cabi.check(lib.ngp_grid_level_scales(16, w["S"], 16, _cabi.ptr(scales), _cabi.stream()))
    c = orc.render(o, d, np.array([-1, -1, -1, 1, 1, 1], np.float32), frame["bitfield"], w["table"], w["offsets"], w["S"], 16, w["w_sigma"],
                   w["w_color"], density_scale=w["density_scale"], scales=scales.cpu().numpy())
    out = render_ours(ours, torch.from_numpy(o).to(cuda), torch.from_numpy(d).to(cuda))
    img, dep = out["image"][0].cpu().numpy(), out["depth"][0].cpu().numpy()
    assert np.abs(img - c["image"]).max() <= 1e-3
    ok = ~np.isnan(c["depth"])
    assert np.array_equal(np.isnan(dep), ~ok) and np.abs(dep[ok] - c["depth"][ok]).max() <= 1e-3
    spr = out["samples_per_ray"].cpu().numpy()
    if variant == "A":
        assert np.array_equal(spr, c["samples_per_ray"]) and out["stats"][0].item() == c["iterations"]
    else:
        assert (spr != c["samples_per_ray"]).sum() <= 8


def test_render_batched_frames_and_edge_cases(frame, cuda):
    """[B,N,3] batches share one n_step schedule (renderer.py:265-269); empty and all-miss inputs."""
    ours = frame["oursB"]
    rays = [syn.pinhole_rays(syn.orbit_pose(az), 64, 64) for az in (0.0, 90.0, 200.0)]
    o = torch.from_numpy(np.stack([r[0] for r in rays])).to(cuda)
    d = torch.from_numpy(np.stack([r[1] for r in rays])).to(cuda)
    with torch.no_grad():
        out = ours.render(o, d, bg_color=1, perturb=False)
        assert out["image"].shape == (3, 4096, 3) and out["depth"].shape == (3, 4096)
        flat = ours.render(o.view(1, -1, 3), d.view(1, -1, 3), bg_color=1, perturb=False)
        assert torch.equal(out["image"].view(-1, 3), flat["image"].view(-1, 3))
        # all rays miss the box: image = bg, depth NaN
        om = torch.full((1, 1000, 3), 5.0, device=cuda)
        dm = torch.tensor([0.0, 1.0, 0.0], device=cuda).expand(1, 1000, 3).contiguous()
        res = ours.render(om, dm, bg_color=1, perturb=False)
        assert torch.equal(res["image"], torch.ones_like(res["image"])) and torch.isnan(res["depth"]).all()
        empty = ours.render(torch.empty(1, 0, 3, device=cuda), torch.empty(1, 0, 3, device=cuda), bg_color=1)
        assert empty["image"].shape == (1, 0, 3)
        # determinism: the persistent kernel's tile order varies, the result must not
        again = ours.render(o, d, bg_color=1, perturb=False)
        assert torch.equal(again["image"], out["image"]) and torch.equal(torch.nan_to_num(again["depth"]), torch.nan_to_num(out["depth"]))
