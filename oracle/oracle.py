"""numpy/ctypes front-end of the CPU oracle (oracle/ngp_oracle.c).

TEST INFRASTRUCTURE ONLY -- may be imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by nerfsafetyvalidation_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "ngp_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libngp_oracle.so")

CFLAGS = ["-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-mf16c", "-mfma",
          "-fvisibility=hidden", "-shared", "-fPIC", "-Wall", "-Wno-unknown-pragmas"]


def build(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.check_call(["gcc", *CFLAGS, SRC, "-o", LIB, "-lm"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.orc_compact_alive.restype = C.c_uint32
        _lib.orc_render.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
    o, d, bb = _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3), _f32(aabb)
    N = o.shape[0]
    nears, fars = np.empty(N, np.float32), np.empty(N, np.float32)
    lib().orc_near_far_from_aabb(_p(o), _p(d), _p(bb), C.c_uint32(N), C.c_float(min_near), _p(nears), _p(fars))
    return nears, fars


def morton3D(coords):
    c = _i32(coords)
    out = np.empty(c.shape[0], np.int32)
    lib().orc_morton3D(_p(c), C.c_uint32(c.shape[0]), _p(out))
    return out


def morton3D_invert(indices):
    i = _i32(indices)
    out = np.empty((i.shape[0], 3), np.int32)
    lib().orc_morton3D_invert(_p(i), C.c_uint32(i.shape[0]), _p(out))
    return out


def packbits(grid, thresh):
    g = _f32(grid)
    n = g.size // 8
    out = np.empty(n, np.uint8)
    lib().orc_packbits(_p(g), C.c_uint32(n), C.c_float(thresh), _p(out))
    return out


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, bitfield, cascade, H, nears, fars,
               align=-1, dt_gamma=0.0, max_steps=1024):
    M = n_alive * n_step
    if align > 0:
        M += align - (M % align)
    xyzs, dirs, deltas = (np.zeros((M, 3), np.float32), np.zeros((M, 3), np.float32), np.zeros((M, 2), np.float32))
    ra, rt, o, d = _i32(rays_alive), _f32(rays_t), _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3)
    bf, ne, fa = np.ascontiguousarray(bitfield, np.uint8), _f32(nears), _f32(fars)
    lib().orc_march_rays(C.c_uint32(n_alive), C.c_uint32(n_step), _p(ra), _p(rt), _p(o), _p(d), C.c_float(bound),
                         C.c_float(dt_gamma), C.c_uint32(max_steps), C.c_uint32(cascade), C.c_uint32(H), _p(bf),
                         _p(ne), _p(fa), _p(xyzs), _p(dirs), _p(deltas))
    return xyzs, dirs, deltas


def composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image):
    """In place on the numpy arrays passed (must be contiguous int32 / float32)."""
    for a in (rays_alive, rays_t, weights_sum, depth, image):
        assert a.flags["C_CONTIGUOUS"]
    s, c, dl = _f32(sigmas), _f32(rgbs), _f32(deltas)
    lib().orc_composite_rays(C.c_uint32(n_alive), C.c_uint32(n_step), _p(rays_alive), _p(rays_t), _p(s), _p(c), _p(dl),
                             _p(weights_sum), _p(depth), _p(image))


def compact_alive(rays_alive):
    a = _i32(rays_alive)
    out = np.empty_like(a)
    k = lib().orc_compact_alive(_p(a), C.c_uint32(a.shape[0]), _p(out))
    return out[:k].copy()


def grid_level_scales(L, S, H):
    out = np.empty(L, np.float32)
    lib().orc_grid_level_scales(C.c_uint32(L), C.c_float(S), C.c_uint32(H), _p(out))
    return out


def grid_encode(inputs01, table, offsets, S, H, scales=None, gridtype=0, align_corners=False,
                round_mode=1, out_layout=1, want_indices=False):
    """table: [rows, C] float32 or float16 (float16 -> half outputs). Returns ([B, L*C] or [L,B,C], idx|None)."""
    x = _f32(inputs01).reshape(-1, 3)
    B = x.shape[0]
    f16 = table.dtype == np.float16
    t = np.ascontiguousarray(table)
    Cc = t.shape[1]
    off = _i32(offsets)
    L = off.shape[0] - 1
    out = np.empty((B, L * Cc) if out_layout == 1 else (L, B, Cc), np.float16 if f16 else np.float32)
    idx = np.empty((L, B, 8), np.int32) if want_indices else None
    sc = None if scales is None else _f32(scales)
    lib().orc_grid_encode(_p(x), _p(t), C.c_int(int(f16)), _p(off), _p(out), C.c_uint32(B), C.c_uint32(Cc),
                          C.c_uint32(L), _p(sc), C.c_float(S), C.c_uint32(H), C.c_uint32(gridtype),
                          C.c_int(int(align_corners)), C.c_int(round_mode), C.c_int(out_layout), _p(idx))
    return out, idx


def sh_encode(dirs, degree=4):
    d = _f32(dirs).reshape(-1, 3)
    out = np.empty((d.shape[0], degree * degree), np.float32)
    lib().orc_sh_encode(_p(d), _p(out), C.c_uint32(d.shape[0]), C.c_uint32(degree))
    return out


def ffmlp(inputs_f16, weights_f16, in_dim, out_dim, hidden, num_layers):
    x = np.ascontiguousarray(inputs_f16, np.float16)
    w = np.ascontiguousarray(weights_f16, np.float16)
    out = np.empty((x.shape[0], out_dim), np.float16)
    lib().orc_ffmlp(_p(x), _p(w), C.c_uint32(x.shape[0]), C.c_uint32(in_dim), C.c_uint32(out_dim), C.c_uint32(hidden),
                    C.c_uint32(num_layers), _p(out))
    return out


def network_forward(xyzs, dirs, table_f16, offsets, S, H, w_sigma, w_color, nl_sigma=2, nl_color=3, bound=1.0,
                    density_scale=1.0, scales=None, grid_round_mode=0):
    x, d = _f32(xyzs).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
    M = x.shape[0]
    t = np.ascontiguousarray(table_f16, np.float16)
    off = _i32(offsets)
    ws, wc = np.ascontiguousarray(w_sigma, np.float16), np.ascontiguousarray(w_color, np.float16)
    sig, rgb = np.empty(M, np.float32), np.empty((M, 3), np.float32)
    sc = None if scales is None else _f32(scales)
    lib().orc_network_forward(_p(x), _p(d), C.c_uint32(M), _p(t), _p(off), C.c_uint32(off.shape[0] - 1), C.c_float(S),
                              C.c_uint32(H), _p(sc), _p(ws), C.c_uint32(nl_sigma), _p(wc), C.c_uint32(nl_color),
                              C.c_float(bound), C.c_float(density_scale), C.c_int(grid_round_mode), _p(sig), _p(rgb))
    return sig, rgb


def render(rays_o, rays_d, aabb, bitfield, table_f16, offsets, S, Hbase, w_sigma, w_color, *, min_near=0.2, bound=1.0,
           dt_gamma=0.0, max_steps=1024, cascade=1, grid_size=128, nl_sigma=2, nl_color=3, density_scale=1.0,
           bg_color=1.0, scales=None, grid_round_mode=0):
    """Whole run_cuda inference loop. Returns dict(image, depth, weights_sum, samples_per_ray, iterations, samples)."""
    o, d = _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3)
    N = o.shape[0]
    bb, bf = _f32(aabb), np.ascontiguousarray(bitfield, np.uint8)
    t, off = np.ascontiguousarray(table_f16, np.float16), _i32(offsets)
    ws, wc = np.ascontiguousarray(w_sigma, np.float16), np.ascontiguousarray(w_color, np.float16)
    image, depth, wsum = np.empty((N, 3), np.float32), np.empty(N, np.float32), np.empty(N, np.float32)
    spr, stats = np.empty(N, np.int32), np.zeros(3, np.int64)
    sc = None if scales is None else _f32(scales)
    rc = lib().orc_render(_p(o), _p(d), C.c_uint32(N), _p(bb), C.c_float(min_near), C.c_float(bound), C.c_float(dt_gamma),
                          C.c_uint32(max_steps), C.c_uint32(cascade), C.c_uint32(grid_size), _p(bf), _p(t), _p(off),
                          C.c_uint32(off.shape[0] - 1), C.c_float(S), C.c_uint32(Hbase), _p(sc), _p(ws),
                          C.c_uint32(nl_sigma), _p(wc), C.c_uint32(nl_color), C.c_float(density_scale),
                          C.c_float(bg_color), C.c_int(grid_round_mode), _p(image), _p(depth), _p(wsum), _p(spr), _p(stats))
    assert rc == 0
    return dict(image=image, depth=depth, weights_sum=wsum, samples_per_ray=spr, iterations=int(stats[0]),
                samples=int(stats[1]), padded_rows=int(stats[2]))


if __name__ == "__main__":
    print(build(force=True))
