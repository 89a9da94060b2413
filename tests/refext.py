"""GPU oracle: the reference's OWN CUDA extensions, rebuilt unmodified for sm_100a by oracle/build_ref.py
(oracle/_ref/_raymarching.so, _gridencoder.so, _shencoder.so, _ffmlp.so), driven by a restatement of the thin Python
glue around them.  TEST INFRASTRUCTURE ONLY.

/root/reference does not exist on the GPU box, so the reference's Python wrappers cannot be imported there; the few
lines of glue they contain are restated here with file:line citations, under the dtype flow the reference has inside
`torch.autocast('cuda', dtype=torch.float16)`:
    raymarching/raymarching.py:19-49,292-362   near_far_from_aabb / march_rays / composite_rays wrappers
    gridencoder/grid.py:19-59,143-156          (x+bound)/(2 bound), half table, [L,B,C] -> [B,L*C]
    shencoder/sphere_harmonics.py:14-41        fp32
    ffmlp/ffmlp.py:15-48,146-169               half, always-pad to 128 rows
    nerf/network_ff.py:51-74                   forward
    nerf/renderer.py:261-284,329-386           run_cuda inference loop
"""
import importlib.util
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
_mods = {}


def available(name="_raymarching"):
    return os.path.exists(os.path.join(REF_DIR, name + ".so"))


def ext(name):
    """Imports oracle/_ref/<name>.so (a pybind11 module built from the reference's sources)."""
    if name not in _mods:
        path = os.path.join(REF_DIR, name + ".so")
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _mods[name] = mod
    return _mods[name]


# ------------------------------------------------------------------ raymarching wrappers
def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
    rays_o = rays_o.contiguous().view(-1, 3)
    rays_d = rays_d.contiguous().view(-1, 3)
    N = rays_o.shape[0]
    nears = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
    fars = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
    ext("_raymarching").near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars)
    return nears, fars


def morton3D(coords):
    N = coords.shape[0]
    indices = torch.empty(N, dtype=torch.int32, device=coords.device)
    ext("_raymarching").morton3D(coords.int().contiguous(), N, indices)
    return indices


def morton3D_invert(indices):
    N = indices.shape[0]
    coords = torch.empty(N, 3, dtype=torch.int32, device=indices.device)
    ext("_raymarching").morton3D_invert(indices.int().contiguous(), N, coords)
    return coords


def packbits(grid, thresh):
    grid = grid.contiguous()
    N = grid.shape[0] * grid.shape[1] // 8
    bitfield = torch.empty(N, dtype=torch.uint8, device=grid.device)
    ext("_raymarching").packbits(grid, N, thresh, bitfield)
    return bitfield


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, density_bitfield, C, H, near, far, align=-1,
               perturb=False, dt_gamma=0, max_steps=1024):
    M = n_alive * n_step
    if align > 0:
        M += align - (M % align)
    xyzs = torch.zeros(M, 3, dtype=rays_o.dtype, device=rays_o.device)
    dirs = torch.zeros(M, 3, dtype=rays_o.dtype, device=rays_o.device)
    deltas = torch.zeros(M, 2, dtype=rays_o.dtype, device=rays_o.device)
    ext("_raymarching").march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, dt_gamma, max_steps, C, H,
                                   density_bitfield, near, far, xyzs, dirs, deltas, perturb)
    return xyzs, dirs, deltas


def composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image):
    ext("_raymarching").composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas.float().contiguous(), rgbs.float().contiguous(),
                                       deltas.float().contiguous(), weights_sum, depth, image)


# ------------------------------------------------------------------ encoders / MLP
def grid_encode(inputs01, embeddings, offsets, per_level_scale, base_resolution=16, gridtype=0, align_corners=False):
    """grid.py:19-59: returns [B, L*C] in embeddings' dtype (pass a half table for the autocast flow)."""
    inputs01 = inputs01.contiguous()
    B, D = inputs01.shape
    L = offsets.shape[0] - 1
    C = embeddings.shape[1]
    S = np.log2(per_level_scale)
    outputs = torch.empty(L, B, C, device=inputs01.device, dtype=embeddings.dtype)
    dy_dx = torch.empty(1, device=inputs01.device, dtype=embeddings.dtype)
    ext("_gridencoder").grid_encode_forward(inputs01, embeddings, offsets, outputs, B, D, C, L, S, base_resolution, False, dy_dx,
                                            gridtype, align_corners)
    return outputs.permute(1, 0, 2).reshape(B, L * C)


def sh_encode(dirs, degree=4):
    dirs = dirs.float().contiguous()
    B = dirs.shape[0]
    outputs = torch.empty(B, degree ** 2, dtype=torch.float32, device=dirs.device)
    dy_dx = torch.empty(1, dtype=torch.float32, device=dirs.device)
    ext("_shencoder").sh_encode_forward(dirs, outputs, B, 3, degree, False, dy_dx)
    return outputs


def ffmlp(inputs, weights_f16, input_dim, output_dim, hidden_dim, num_layers):
    """FFMLP.forward (ffmlp.py:146-169) in eval mode: pad to 128 rows (always adds rows), ffmlp_inference, slice."""
    B, C = inputs.shape
    inputs = inputs.to(torch.half)
    pad = 128 - (B % 128)
    inputs = torch.cat([inputs, torch.zeros(pad, C, dtype=inputs.dtype, device=inputs.device)], dim=0).contiguous()
    padded_out = 16
    outputs = torch.empty(inputs.shape[0], padded_out, device=inputs.device, dtype=torch.half)
    buf = torch.empty(inputs.shape[0], hidden_dim, device=inputs.device, dtype=torch.half)
    ext("_ffmlp").ffmlp_inference(inputs, weights_f16.contiguous(), inputs.shape[0], input_dim, padded_out, hidden_dim, num_layers, 0, 6,
                                  buf, outputs)
    return outputs[:B, :output_dim]


def reference_renderer_class():
    """The reference's OWN nerf/renderer.py::NeRFRenderer (copied unmodified to oracle/_ref/pyref by oracle/build_ref.py),
    imported behind three stubs: `raymarching` = the wrappers above on the rebuilt extension, `trimesh` (export_mesh only) and
    `nerf.utils.custom_meshgrid`.  None when the copy is not there."""
    import sys
    import types
    path = os.path.join(REF_DIR, "pyref", "nerf", "renderer.py")
    if not os.path.exists(path) or not available("_raymarching"):
        return None
    if "ngp_ref_nerf.renderer" in sys.modules:
        return sys.modules["ngp_ref_nerf.renderer"].NeRFRenderer
    rm = types.ModuleType("raymarching")
    rm.near_far_from_aabb, rm.morton3D, rm.morton3D_invert = near_far_from_aabb, morton3D, morton3D_invert
    rm.march_rays, rm.composite_rays = march_rays, composite_rays
    rm.packbits = lambda grid, thresh, bitfield=None: packbits(grid, thresh)   # raymarching.py:129-155 (allocates when None)
    utils = types.ModuleType("ngp_ref_nerf.utils")
    utils.custom_meshgrid = lambda *a: torch.meshgrid(*a, indexing="ij")       # nerf/utils.py:28-33
    pkg = types.ModuleType("ngp_ref_nerf")
    pkg.__path__ = [os.path.dirname(path)]
    saved = {k: sys.modules.get(k) for k in ("raymarching", "trimesh")}
    sys.modules.update({"raymarching": rm, "trimesh": saved["trimesh"] or types.ModuleType("trimesh"), "ngp_ref_nerf": pkg,
                        "ngp_ref_nerf.utils": utils})
    try:
        spec = importlib.util.spec_from_file_location("ngp_ref_nerf.renderer", path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules["ngp_ref_nerf.renderer"] = mod
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():   # the stubs stay bound inside the imported module; the process-wide names are restored
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod.NeRFRenderer


def reference_training_stack():
    """The reference's OWN Python for the differentiable pieces, imported unmodified from oracle/_ref/pyref on top of the rebuilt
    extensions: returns (a constructor of nerf/network_ff.py's NeRFNetwork on the reference's gridencoder / shencoder / ffmlp packages and its
    own nerf/renderer.py, the reference's `raymarching` module with march_rays_train / composite_rays_train).  The packages'
    `try: import _<ext> as _backend` (e.g. gridencoder/grid.py:9-12) finds the rebuilt .so files through sys.modules.  Everything is
    imported under private module names and the process-wide names are restored, so the product's packages are not shadowed.
    None when the copies or an extension are missing."""
    import sys
    import types
    root = os.path.join(REF_DIR, "pyref")
    if "ngp_ref_train" in _mods:
        return _mods["ngp_ref_train"]
    if not all(available(n) for n in ("_raymarching", "_gridencoder", "_shencoder", "_ffmlp")) or \
            not os.path.exists(os.path.join(root, "gridencoder", "grid.py")):
        return None
    names = ["_raymarching", "_gridencoder", "_shencoder", "_ffmlp", "raymarching", "gridencoder", "shencoder", "ffmlp", "encoding",
             "activation", "nerf", "nerf.renderer", "nerf.network_ff", "nerf.utils", "trimesh", "turtle"]
    subs = ["raymarching.raymarching", "gridencoder.grid", "shencoder.sphere_harmonics", "ffmlp.ffmlp"]
    saved = {k: sys.modules.get(k) for k in names + subs}
    for k in names + subs:
        sys.modules.pop(k, None)
    for n in ("_raymarching", "_gridencoder", "_shencoder", "_ffmlp"):
        sys.modules[n] = ext(n)
    sys.modules["trimesh"] = saved["trimesh"] or types.ModuleType("trimesh")
    if "turtle" not in sys.modules:   # ffmlp/ffmlp.py:2 carries a stray `from turtle import backward, forward` (needs tkinter); unused
        turtle = types.ModuleType("turtle")
        turtle.backward = turtle.forward = None
        sys.modules["turtle"] = turtle
    utils = types.ModuleType("nerf.utils")
    utils.custom_meshgrid = lambda *a: torch.meshgrid(*a, indexing="ij")       # nerf/utils.py:28-33
    pkg = types.ModuleType("nerf")
    pkg.__path__ = [os.path.join(root, "nerf")]
    sys.modules["nerf"], sys.modules["nerf.utils"] = pkg, utils
    sys.path.insert(0, root)
    try:
        import importlib
        rm = importlib.import_module("raymarching")
        net = importlib.import_module("nerf.network_ff")
        lazy = {k: importlib.import_module(k) for k in ("gridencoder", "shencoder", "ffmlp")}   # encoding.py:60-70 imports them per call

        def make(**kwargs):
            """NeRFNetwork(**kwargs) with the reference's encoder packages visible to encoding.get_encoder's call-time imports."""
            held = {k: sys.modules.get(k) for k in lazy}
            sys.modules.update(lazy)
            try:
                return net.NeRFNetwork(**kwargs)
            finally:
                for k, v in held.items():
                    if v is None:
                        sys.modules.pop(k, None)
                    else:
                        sys.modules[k] = v
        out = (make, rm)
    finally:
        sys.path.remove(root)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    _mods["ngp_ref_train"] = out
    return out


class RefModel:
    """Holds what network_ff.NeRFNetwork holds, as device tensors (half copies = what autocast would produce)."""

    def __init__(self, weights, bitfield, device="cuda", cascade=1, grid_size=128, min_near=0.2):
        self.table = torch.from_numpy(weights["table"]).to(device)
        self.offsets = torch.from_numpy(weights["offsets"]).to(device)
        self.per_level_scale = weights["per_level_scale"]
        self.w_sigma = torch.from_numpy(weights["w_sigma"]).to(device)
        self.w_color = torch.from_numpy(weights["w_color"]).to(device)
        self.num_layers, self.num_layers_color = weights["num_layers"], weights["num_layers_color"]
        self.bound, self.density_scale = weights["bound"], weights["density_scale"]
        self.bitfield = torch.from_numpy(np.ascontiguousarray(bitfield)).to(device)
        self.cascade, self.grid_size, self.min_near = cascade, grid_size, min_near
        b = self.bound
        self.aabb = torch.tensor([-b, -b, -b, b, b, b], dtype=torch.float32, device=device)

    def density(self, x):
        """network_ff.py:76-89 under autocast: {'sigma': exp(h0) fp32} (geo_feat is not needed by the occupancy update)."""
        x01 = (x + self.bound) / (2 * self.bound)
        h = ffmlp(grid_encode(x01, self.table, self.offsets, self.per_level_scale), self.w_sigma, 32, 16, 64, self.num_layers)
        return {"sigma": torch.exp(h[..., 0].float())}

    def forward(self, x, d):
        """network_ff.py:51-74 under autocast."""
        x01 = (x + self.bound) / (2 * self.bound)
        enc = grid_encode(x01, self.table, self.offsets, self.per_level_scale)
        h = ffmlp(enc, self.w_sigma, 32, 16, 64, self.num_layers)
        sigma = torch.exp(h[..., 0].float())
        geo_feat = h[..., 1:]
        dsh = sh_encode(d)
        p = torch.zeros_like(geo_feat[..., :1])
        hin = torch.cat([dsh, geo_feat.float(), p.float()], dim=-1)   # autocast promotes cat to fp32; FFMLP casts to half
        hc = ffmlp(hin, self.w_color, 32, 3, 64, self.num_layers_color)
        rgb = torch.sigmoid(hc)   # half in, half out
        return sigma, rgb

    @torch.no_grad()
    def run_cuda(self, rays_o, rays_d, dt_gamma=0, bg_color=1, max_steps=1024, record=None, count=True):
        """renderer.py:261-284,329-386.  `record`: optional list receiving per-iteration snapshots (teacher forcing).
        `count=False` runs the loop exactly as the reference does (no per-ray sample counting, no extra host sync): the form
        bench.py times as `reference_gpu`."""
        rays_o = rays_o.contiguous().view(-1, 3)
        rays_d = rays_d.contiguous().view(-1, 3)
        N = rays_o.shape[0]
        device = rays_o.device
        nears, fars = near_far_from_aabb(rays_o, rays_d, self.aabb, self.min_near)
        weights_sum = torch.zeros(N, dtype=torch.float32, device=device)
        depth = torch.zeros(N, dtype=torch.float32, device=device)
        image = torch.zeros(N, 3, dtype=torch.float32, device=device)
        spr = torch.zeros(N, dtype=torch.int32, device=device)
        n_alive = N
        rays_alive = torch.arange(n_alive, dtype=torch.int32, device=device)
        rays_t = nears.clone()
        step = iters = samples = 0
        sigmas = rgbs = None
        while step < max_steps:
            n_alive = rays_alive.shape[0]
            if n_alive <= 0:
                break
            n_step = max(min(N // n_alive, 8), 1)
            xyzs, dirs, deltas = march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, self.bound, self.bitfield,
                                            self.cascade, self.grid_size, nears, fars, 128, False, dt_gamma, max_steps)
            sigmas, rgbs = self.forward(xyzs, dirs)
            sigmas = self.density_scale * sigmas
            if record is not None:
                record.append(dict(n_alive=n_alive, n_step=n_step, rays_alive=rays_alive.clone(), rays_t=rays_t.clone(),
                                   xyzs=xyzs, dirs=dirs, deltas=deltas, sigmas=sigmas.clone(), rgbs=rgbs.clone(),
                                   weights_sum=weights_sum.clone(), depth=depth.clone(), image=image.clone()))
            if count:
                cnt = (deltas[: n_alive * n_step, 0] != 0).view(n_alive, n_step).sum(1).int()
                spr.index_add_(0, rays_alive.long(), cnt)
                samples += int(cnt.sum().item())
            composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image)
            if record is not None:
                record[-1].update(alive_after=rays_alive.clone(), rays_t_after=rays_t.clone(), weights_sum_after=weights_sum.clone(),
                                  depth_after=depth.clone(), image_after=image.clone())
            rays_alive = rays_alive[rays_alive >= 0]
            step += n_step
            iters += 1
        image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
        depth = torch.clamp(depth - nears, min=0) / (fars - nears)
        return dict(image=image, depth=depth, weights_sum=weights_sum, samples_per_ray=spr if count else None, iterations=iters,
                    samples=samples if count else None, nears=nears, fars=fars, sigmas=sigmas, rgbs=rgbs)
