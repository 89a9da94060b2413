"""Generates tests/golden/*.npz from the reference's OWN CUDA extensions rebuilt for sm_100a (oracle/_ref/*.so).

Run on a B200 box (the reference kernels need a GPU):
    python tests/golden/make_golden.py [out_dir]          (default: gpurun_out/golden; copy into tests/golden/)

Every input is regenerated from seeds by tests/inputs.py and nerfsafetyvalidation_b200/synthetic.py, so the fixtures
hold mostly OUTPUTS of the reference kernels (plus the few inputs that come out of the reference loop itself).  The CPU
oracle (oracle/ngp_oracle.c) is pinned against these files in tests/test_oracle_golden.py; the CUDA product is checked
against the same files in tests/test_gpu_golden.py.
"""
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from nerfsafetyvalidation_b200 import synthetic as syn  # noqa: E402
from tests import refext  # noqa: E402
from tests.inputs import grid_inputs, make_rays, unit_dirs  # noqa: E402

GOLDEN_FRAME = dict(H=48, W=48, azimuth=110.0)     # the e2e / march / composite fixture frame
GRID_LOG2T = 14                                    # small table: levels 0-2 dense, 3-15 hashed
N_RAYS, N_GRID, N_SH, N_MLP = 4096, 4096, 1024, 512


def np_(t):
    return t.detach().cpu().numpy()


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    dev = torch.device("cuda:0")
    meta = dict(torch=torch.__version__, gpu=torch.cuda.get_device_name(0),
                nvcc=subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.strip().splitlines()[-1])

    # ---- near/far, morton, packbits
    o, d = make_rays(N_RAYS, 0, dev)
    out = {}
    for tag, b, mn in (("b1", 1.0, 0.2), ("b2", 2.0, 0.05)):
        aabb = torch.tensor([-b] * 3 + [b] * 3, dtype=torch.float32, device=dev)
        n, f = refext.near_far_from_aabb(o, d, aabb, mn)
        out["nears_" + tag], out["fars_" + tag] = np_(n), np_(f)
    g = torch.Generator().manual_seed(21)
    coords = torch.randint(0, 1024, (N_RAYS, 3), generator=g, dtype=torch.int32).to(dev)
    out["morton_coords"] = np_(coords)
    out["morton_idx"] = np_(refext.morton3D(coords))
    idx = torch.randint(0, 2 ** 30, (N_RAYS,), generator=g, dtype=torch.int32).to(dev)
    out["invert_idx"] = np_(idx)
    out["invert_coords"] = np_(refext.morton3D_invert(idx))
    grid = torch.rand(1, 65536, generator=g)
    grid[0, ::5] = 0.5
    out["packbits_grid_seed"] = np.int64(21)
    out["packbits_0p5"] = np_(refext.packbits(grid.to(dev), 0.5))
    np.savez_compressed(os.path.join(out_dir, "raymarch_setup.npz"), **out)

    # ---- march / composite snapshots + e2e from the reference loop (variants A and B)
    bitfield, _ = syn.stonehenge_bitfield()
    fo, fd = syn.pinhole_rays(syn.orbit_pose(GOLDEN_FRAME["azimuth"]), GOLDEN_FRAME["H"], GOLDEN_FRAME["W"])
    fo_t, fd_t = torch.from_numpy(fo).to(dev), torch.from_numpy(fd).to(dev)
    for var in "AB":
        w = syn.ngp_weights(0, variant=var)
        ref = refext.RefModel(w, bitfield, device=dev)
        rec = []
        r = ref.run_cuda(fo_t, fd_t, record=rec)
        out = dict(image=np_(r["image"]), depth=np_(r["depth"]), weights_sum=np_(r["weights_sum"]), samples_per_ray=np_(r["samples_per_ray"]),
                   iterations=np.int64(r["iterations"]), samples=np.int64(r["samples"]), nears=np_(r["nears"]), fars=np_(r["fars"]))
        picks = sorted({0, 1, 5, len(rec) // 2, len(rec) - 1})
        out["picks"] = np.asarray(picks, np.int64)
        for i in picks:
            s = rec[i]
            n = s["n_alive"] * s["n_step"]
            out[f"it{i}_n_step"] = np.int64(s["n_step"])
            for k in ("rays_alive", "rays_t", "alive_after", "rays_t_after", "weights_sum", "depth", "image", "weights_sum_after",
                      "depth_after", "image_after"):
                out[f"it{i}_{k}"] = np_(s[k])
            for k in ("xyzs", "dirs", "deltas", "sigmas", "rgbs"):
                out[f"it{i}_{k}"] = np_(s[k][:n].float())
            out[f"it{i}_M"] = np.int64(s["xyzs"].shape[0])
        np.savez_compressed(os.path.join(out_dir, f"render_{var}.npz"), **out)

    # ---- march on random rays, cascade 2, exponential stepping
    bound, cascade, dt_gamma, n_step = 2.0, 2, 1.0 / 128, 4
    o2 = o * bound / 2
    aabb = torch.tensor([-bound] * 3 + [bound] * 3, dtype=torch.float32, device=dev)
    nears, fars = refext.near_far_from_aabb(o2, d, aabb, 0.05)
    g = torch.Generator().manual_seed(11)
    bf = torch.randint(0, 256, (cascade * 128 ** 3 // 8,), generator=g, dtype=torch.uint8)
    alive = torch.randperm(N_RAYS, generator=g)[:3000].int()
    rays_t = nears + torch.rand(N_RAYS, generator=g).to(dev) * 0.3
    x, dd, dl = refext.march_rays(3000, n_step, alive.to(dev), rays_t, o2, d, bound, bf.to(dev), cascade, 128, nears, fars, 128, False,
                                  dt_gamma, 1024)
    np.savez_compressed(os.path.join(out_dir, "march_random.npz"), xyzs=np_(x), dirs=np_(dd), deltas=np_(dl), rays_t=np_(rays_t),
                        nears=np_(nears), fars=np_(fars), alive=alive.numpy())

    # ---- grid encoder (fp16 + fp32 tables), SH, FFMLP
    w = syn.ngp_weights(0, log2_hashmap_size=GRID_LOG2T, variant="B")
    x = grid_inputs(N_GRID, 5, dev)
    offs = torch.from_numpy(w["offsets"]).to(dev)
    out = {}
    for key, tag in (("table", "f16"), ("table_f32", "f32")):
        y = refext.grid_encode(x, torch.from_numpy(w[key]).to(dev), offs, w["per_level_scale"])
        out["grid_" + tag] = np_(y)
    # the per-level scale as the DEVICE evaluates exp2f (ex2.approx), needed by any CPU restatement
    lv = torch.arange(16, dtype=torch.float32, device=dev)
    out["scales"] = np_(torch.exp2(lv * np.float32(w["S"])) * 16 - 1)
    dirs = unit_dirs(N_SH, 2, dev)
    dirs[N_SH // 2:] *= 1.7
    out["sh4"] = np_(refext.sh_encode(dirs, 4))
    w19 = syn.ngp_weights(0)
    g = torch.Generator().manual_seed(34)
    xin = (torch.randn(N_MLP, 32, generator=g) * 0.5).half().to(dev)
    out["mlp_sigma"] = np_(refext.ffmlp(xin, torch.from_numpy(w19["w_sigma"]).to(dev), 32, 16, 64, 2))
    out["mlp_color"] = np_(refext.ffmlp(xin, torch.from_numpy(w19["w_color"]).to(dev), 32, 16, 64, 3))
    np.savez_compressed(os.path.join(out_dir, "encoders_mlp.npz"), **out)

    with open(os.path.join(out_dir, "META.txt"), "w") as f:
        for k, v in meta.items():
            f.write(f"{k}: {v}\n")
        f.write("source: reference extensions rebuilt unmodified for sm_100a by oracle/build_ref.py\n")
    print("golden vectors written to", out_dir)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
