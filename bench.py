#!/usr/bin/env python
"""bench.py -- rendered Mrays/s of the NGP `cuda_ray` inference path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2|cfg5|rollout]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic rays:
    cfg2 (default)  one 800x800 frame of the Stonehenge-shaped scene, 16-level 2^19 hash grid + SH + 64-wide MLPs,
                    variant-A weights (reference init), max_steps 1024, dt_gamma 0, bg 1   (BASELINE.json configs[1])
Every rank renders its own camera of the same orbit (disturbance rollouts / camera batches shard with replicated
weights, no data-path collective) => weak scaling; `value` = rays of all ranks / max-over-ranks time.

`value`    device-resident inputs, CUDA events around ngp_render on the launching stream, L2 flushed between steps.
`e2e`      the same frame through ngp_render_host: pinned host rays -> H2D -> render -> D2H image+depth, inside the timer.
`roofline` dominant kernel = k_render (99.9 % of a step, profiles/): algorithmic 512 B gathered per sample (16 levels x 8
           corners x 2 ch x 2 B) x samples counted by the kernel, over the measured HBM copy peak
           (MEASURED_PEAKS.json).  `traffic` = DRAM bytes of one launch from the committed ncu capture of THIS build and workload
           (profiles/r02_ncu_metrics.json).  The 2^19 table (23.3 MB fp16) is L2-resident on B200, so beside the HBM figure the
           line carries the roof that fits the access pattern -- `l1_wavefront`: ncu-counted L1 wavefronts of the gather at one
           per cycle per SM -- and the tensor-pipe fraction.
`cpu_baseline`  oracle/ngp_oracle.c (OpenMP port of the reference path) on a strided sub-sample of the same frame.
`--impl reference`  the same CPU port on all host threads, as the reference arm (the reference has no CPU
           implementation of the NGP path; its CUDA extensions rebuilt for sm_100a are timed beside it as `reference_gpu`).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (H, W, focal, log2T, dense_bitfield)
    "cfg2": dict(H=800, W=800, focal=1111.11, log2T=19, dense=False,
                 desc="cfg2: 800x800 Stonehenge-shaped frame, NGP L16 T2^19 F2 + SH4 + 64-wide ffmlp, bound 1, max_steps 1024"),
    "cfg5": dict(H=1080, W=1920, focal=1500.0, log2T=22, dense=True,
                 desc="cfg5: 1920x1080 frame, T2^22 table, dense occupancy, max_steps 1024"),
    "rollout": dict(H=200, W=200, focal=277.78, log2T=19, dense=False, frames=64,
                    desc="cfg3 batch: 64 rollout frames of 200x200 per render call"),
}


# The build the committed ncu metrics (profiles/r02_ncu_metrics.json, made by tools/ncu_metrics_json.py from `ncu --set full` captures of
# k_render on this workload) belong to.  Bump it together with the file; a mismatch reports `traffic: null` instead of a stale number.
PROFILE_BUILD = "r02j"


def ncu_metrics(workload):
    """{dram_bytes, l1_global_load_wavefronts, ...} of one k_render launch of `workload`, or None when no capture of this build exists."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")))
        return d.get(workload) if d.get("build") == PROFILE_BUILD else None
    except Exception:
        return None


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--workload", default="cfg2", choices=list(WORKLOADS))
    p.add_argument("--variant", default="A", choices=["A", "B"])
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-ref-gpu", action="store_true")
    p.add_argument("--cpu-rays", type=int, default=0, help="rays in the CPU sample (0 = auto)")
    p.add_argument("--split-rows", action="store_true",
                   help="single-frame workloads under torchrun: ONE frame, rank r renders the r-th block of image rows (strong scaling; "
                        "SURVEY section 8d cfg5) instead of one frame per rank")
    p.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE",
                   help="ngp_ctx_set_option tunable of the renderer's epoch policy (e.g. rows_per_ray=64); recorded in config.options")
    return p.parse_args()


def frame_rays(wl, rank, n_frames=1, with_poses=False):
    """rays of rank `rank`'s step.  Single frames: cameras on the cfg2 orbit, azimuth advanced per rank.  Rollout batches
    (cfg3): the first n_frames frames of this rank's disturbance rollouts (rollouts.py: drone dynamics + per-rollout seeded
    noise + the simulator's pose conversion), i.e. what NerfSimulator would hand to render() during a Monte Carlo run.
    with_poses: also return the [F,4,4] cam2world poses the rays were generated from."""
    from nerfsafetyvalidation_b200 import rollouts as ro, synthetic as syn
    if n_frames > 1:
        steps = wl.get("rollout_steps", 20)
        ids = ro.shard_ids(10 ** 6, rank, max(int(os.environ.get("WORLD_SIZE", "1")), 1))[: -(-n_frames // steps)]
        o, d, states, _ = ro.rollout_rays(ids, steps, wl["H"], wl["W"], wl["focal"])
        o, d = np.ascontiguousarray(o[:n_frames].reshape(-1, 3)), np.ascontiguousarray(d[:n_frames].reshape(-1, 3))
        poses = np.ascontiguousarray(ro.states_to_ngp_poses(states.reshape(-1, 12))[:n_frames], np.float32)
    else:
        pose = syn.orbit_pose(37.0 + 45.0 * rank)
        o, d = syn.pinhole_rays(pose, wl["H"], wl["W"], focal=wl["focal"])
        poses = np.asarray(pose, np.float32)[None].copy()
    return (o, d, poses) if with_poses else (o, d)


def scene_for(wl):
    from nerfsafetyvalidation_b200 import synthetic as syn
    bitfield, grid = syn.stonehenge_bitfield()
    if wl["dense"]:
        bitfield = np.full_like(bitfield, 255)
    return bitfield, grid


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.thread = threading.Thread(target=self.run, daemon=True)

    def sample(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                 capture_output=True, text=True, timeout=5).stdout.strip()
            if out:
                self.rows.append([c.strip() for c in out.splitlines()[0].split(",")])
        except Exception:
            pass

    def run(self):
        while not self.stop.is_set():
            self.sample()
            self.stop.wait(0.1)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=10)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def cpu_port_mrays(wl, variant, n_rays, threads=None):
    """oracle/ngp_oracle.c on a strided sub-sample of rank 0's frame.  Returns (Mrays/s, seconds, rays, threads)."""
    from nerfsafetyvalidation_b200 import synthetic as syn
    from oracle import oracle as orc
    if threads:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    bitfield, _ = scene_for(wl)
    w = syn.ngp_weights(0, log2_hashmap_size=wl["log2T"], variant=variant)
    o, d = frame_rays(wl, 0)
    stride = max(1, o.shape[0] // n_rays)
    o, d = np.ascontiguousarray(o[::stride]), np.ascontiguousarray(d[::stride])
    aabb = np.array([-1, -1, -1, 1, 1, 1], np.float32)
    t0 = time.perf_counter()
    r = orc.render(o, d, aabb, bitfield, w["table"], w["offsets"], w["S"], 16, w["w_sigma"], w["w_color"],
                   density_scale=w["density_scale"])
    dt = time.perf_counter() - t0
    return o.shape[0] / dt / 1e6, dt, o.shape[0], r


def cpu_run_sampler_mrays(H=100, W=100):
    """BASELINE.json configs[0] on the host cores: vanilla NeRF (frequency encoders, 11 264 params, torch.manual_seed(0) init)
    rendering one 100x100 frame with the reference's pure-PyTorch `run()` sampler (64 + 64 samples, max_ray_batch 4096), restated
    in oracle/run_sampler.py and pinned against the imported reference (tests/golden/pyref_cfg1.npz)."""
    import math

    import torch

    from oracle import run_sampler as rs
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    fl = 0.5 * W / math.tan(0.5 * 0.6911112)
    pose = np.eye(4, dtype=np.float32)
    pose[2, 3] = -2.5
    o, d = rs.get_rays(pose[None], fl, fl, W / 2, H / 2, H, W)
    net = rs.VanillaNeRF(seed=0)
    o, d = torch.from_numpy(o[0]), torch.from_numpy(d[0])
    rs.render(net, o[:512], d[:512])   # warm-up (thread pool, allocator)
    t0 = time.perf_counter()
    out = rs.render(net, o, d, max_ray_batch=4096, num_steps=64, upsample_steps=64, bg_color=1.0)
    dt = time.perf_counter() - t0
    return {"value": H * W / dt / 1e6, "unit": "Mrays/s", "cores": cores, "seconds": dt, "kind": "port",
            "what": f"cfg1: vanilla NeRF (frequency encoder, cuda_ray=False, 64+64 samples) rendering one {H}x{W} frame, "
                    "oracle/run_sampler.py = the reference's NeRFRenderer.run restated in torch fp32 on CPU",
            "mean_weights_sum": float(out["weights_sum"].mean())}


def run_reference_arm(args, wl):
    """--impl reference: the CPU port of the reference path on all host threads, same metric / unit / config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)   # torchrun exports OMP_NUM_THREADS=1; this arm is meant to use every host core
    n = args.cpu_rays or 20000
    vals, secs = [], []
    for i in range(args.warmup + args.steps):
        v, dt, rays, _ = cpu_port_mrays(wl, args.variant, n)
        if i >= args.warmup:
            vals.append(v)
            secs.append(dt)
        if sum(secs) > 150:   # keep the whole run within a few minutes
            break
    value = rays * len(secs) / sum(secs) / 1e6
    line = {"metric": "rendered Mrays/s (NGP hashgrid, cuda_ray inference)", "value": value, "unit": "Mrays/s", "impl": "reference",
            "n_gpus": args.gpus, "steps": len(secs), "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16 storage / f32 accumulate", "data": "synthetic",
            "config": {"workload": wl["desc"], "variant": args.variant},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "port",
                             "sample": f"{rays} rays (every {max(1, wl['H'] * wl['W'] // n)}th ray of the frame) per step, oracle/ngp_oracle.c with OpenMP"},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    try:
        line["cpu_reference_run"] = cpu_run_sampler_mrays()
    except Exception as e:   # reported beside the arm, never instead of it
        line["cpu_reference_run"] = {"unavailable": f"{type(e).__name__}: {e}"}
    print(json.dumps(line))
    return 0


def main():
    args = parse()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference_arm(args, wl)

    # stdout carries exactly ONE line (the JSON): anything libraries print there (NCCL's version banner under NCCL_DEBUG, ...)
    # goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    from nerfsafetyvalidation_b200 import _cabi, synthetic as syn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU fallback (use --impl reference for the CPU port)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- model: rank 0 builds the weights, everyone else receives them with ONE broadcast (replicated weights)
    bitfield, grid = scene_for(wl)
    w = syn.ngp_weights(0 if rank == 0 else 1, log2_hashmap_size=wl["log2T"], variant=args.variant)
    model = syn.build_model(w, bitfield, grid, device=dev)
    if world > 1:
        for t in (model.encoder.embeddings.data, model.sigma_net.weights.data, model.color_net.weights.data, model.density_bitfield):
            dist.broadcast(t, src=0)
        model.invalidate_caches()
    lib = _cabi.load()
    options = {k: int(v) for k, v in (o.split("=", 1) for o in args.opt)}
    for k, v in options.items():
        model._native.set_option(k, v)

    n_frames = wl.get("frames", 1)
    split_rows = args.split_rows and world > 1 and n_frames == 1
    o_np, d_np, poses_np = frame_rays(wl, 0 if split_rows else rank, n_frames, with_poses=True)
    if split_rows:   # contiguous row blocks keep the live-list order within a rank; the n_step schedule becomes per rank
        r0, r1 = wl["H"] * rank // world, wl["H"] * (rank + 1) // world
        o_np, d_np = np.ascontiguousarray(o_np[r0 * wl["W"]:r1 * wl["W"]]), np.ascontiguousarray(d_np[r0 * wl["W"]:r1 * wl["W"]])
    N = o_np.shape[0]
    o_dev, d_dev = torch.from_numpy(o_np).to(dev), torch.from_numpy(d_np).to(dev)
    o_pin, d_pin = torch.from_numpy(o_np).pin_memory(), torch.from_numpy(d_np).pin_memory()
    img_pin = torch.empty(N, 3, dtype=torch.float32).pin_memory()
    dep_pin = torch.empty(N, dtype=torch.float32).pin_memory()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def render_dev(stats=False):
        with torch.no_grad():
            return model.render(o_dev.view(1, N, 3), d_dev.view(1, N, 3), bg_color=1, perturb=False, dt_gamma=0, max_steps=1024,
                                return_stats=stats)

    def render_host():
        model._configure_native(model._native)
        h = model._native.get()
        model._set_scene(model._native, h)
        _cabi.check(lib.ngp_render_host(h, _cabi.ptr(o_pin), _cabi.ptr(d_pin), N, 0.0, 1024, 1.0, _cabi.ptr(img_pin), _cabi.ptr(dep_pin),
                                        _cabi.stream()), "ngp_render_host")

    # the same frames from their 4x4 poses (64 B per frame up instead of 24 B per ray); a row block has no pose of its own
    pose_pin = None if split_rows else torch.from_numpy(poses_np).pin_memory()

    def render_pose_host():
        model.render_poses_host(pose_pin, (wl["focal"], wl["focal"], wl["W"] / 2, wl["H"] / 2), wl["H"], wl["W"], image=img_pin, depth=dep_pin,
                                dt_gamma=0, bg_color=1, max_steps=1024)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, host_timer=False):
        """K steps, L2 flushed before each, each step timed on the device (events) or on the host (sync call)."""
        total = 0.0
        for _ in range(steps):
            flush.zero_()
            if host_timer:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                fn()
                total += (time.perf_counter() - t0) * 1e3
            else:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                e1.synchronize()
                total += e0.elapsed_time(e1)
        return total

    # ---- warm-up (also counts samples / iterations once)
    out = render_dev(stats=True)
    torch.cuda.synchronize()
    stats = out["stats"].cpu().tolist()
    iters, samples = int(stats[0]), int(stats[1])
    epochs, replays = int(stats[2]), int(stats[3])
    assert epochs >= 0, "ngp_render: the kernel's schedule check failed (stats[2] = -1)"
    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        render_dev()
        render_host()

    # ---- timed: device-resident
    barrier()
    launches0 = _cabi.launch_count()
    with ClockSampler(local) as clocks:
        ms_dev = timed(render_dev, args.steps)
        barrier()
        launches = _cabi.launch_count() - launches0
        ms_e2e = timed(render_host, args.steps, host_timer=True)
        barrier()
        ms_pose = None
        if pose_pin is not None:
            render_pose_host()
            ms_pose = timed(render_pose_host, args.steps, host_timer=True)
            barrier()
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e, ms_pose or 0.0], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e, ms_pose = t.tolist()
        ms_pose = ms_pose or None
        cnt = torch.tensor([N, samples], dtype=torch.int64, device=dev)
        dist.all_reduce(cnt)
        rays_all, samples_all = cnt.tolist()
    else:
        rays_all, samples_all = N, samples

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        ms_step = ms_dev / args.steps
        value = rays_all / (ms_step * 1e-3) / 1e6
        e2e_value = rays_all / (ms_e2e / args.steps * 1e-3) / 1e6
        achieved = samples * 512 / (ms_step * 1e-3) / 1e9    # this rank's kernel: its own samples / its own launch time
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        n_sm = torch.cuda.get_device_properties(dev).multi_processor_count
        ncu = None if args.variant != "A" and args.workload == "cfg2" else ncu_metrics(args.workload)
        tensor_peak = float(peaks.get("bf16_tflops_sustained", 1363.0))
        tensor_ach = samples * 36864 / (ms_step * 1e-3) / 1e12
        roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                    "traffic": int(ncu["dram_bytes"]) if ncu else None, "kernel": "k_render", "algorithmic_bytes_per_sample": 512,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                    # The table is served from L2 / L1 (traffic above = what reaches DRAM), so the roof that fits this access pattern is
                    # the L1 data pipe: one wavefront per cycle per SM.  The wavefront count is ncu's, of this build on this workload
                    # (K consecutive samples of a ray sit in adjacent lanes, so coarse levels coalesce); floor = wavefronts / (SMs x clock).
                    "l1_wavefront": ({"wavefronts_per_launch": int(ncu["l1_global_load_wavefronts"]),
                                      "floor_ms": ncu["l1_global_load_wavefronts"] / n_sm / (sm_max * 1e3),
                                      "frac": ncu["l1_global_load_wavefronts"] / n_sm / (sm_max * 1e3) / ms_step,
                                      "source": f"profiles/r02_ncu_metrics.json (build {PROFILE_BUILD}, {ncu['report']})"} if ncu else None),
                    "tensor": {"achieved_tflops": tensor_ach, "peak_tflops": tensor_peak, "frac": tensor_ach / tensor_peak,
                               "what": "36864 algorithmic FLOP per sample (7 layers, 64-wide) over bf16_tflops_sustained; the kernel issues "
                                       "every layer once per 32-row lane quadrant (masked M=128 MMAs), 4x that on the pipe"}}
        line = {"metric": "rendered Mrays/s (NGP hashgrid, cuda_ray inference)", "value": value, "unit": "Mrays/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if split_rows else "weak",
                "vs_baseline": None, "dtype": "f16 storage / f32 accumulate", "data": "synthetic",
                "config": {"workload": wl["desc"], "variant": args.variant, "rays_per_gpu_per_step": N, "samples_per_step": samples,
                           "samples_per_ray": samples / N, "iterations_per_step": iters, "epochs_per_step": epochs, "replayed_epochs": replays, **({"options": options} if options else {}),
                           "frames_per_step": n_frames,
                           "l2": "flushed between steps (256 MB write)", "parallelism": f"replicated weights, {world} ray shards" + (" (row blocks of ONE frame)" if split_rows else "")},
                "frames_per_s": n_frames * (1 if split_rows else world) / (ms_step * 1e-3), "msamples_per_s": samples_all / (ms_step * 1e-3) / 1e6,
                "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(o_pin.numel() * 4 + d_pin.numel() * 4),
                        "d2h_bytes_per_step": int(img_pin.numel() * 4 + dep_pin.numel() * 4), "ms_per_step": ms_e2e / args.steps,
                        "api": "ngp_render_host (C ABI, pinned host buffers)"},
                "e2e_from_pose": None if not ms_pose else {
                    "value": rays_all / (ms_pose / args.steps * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(pose_pin.numel() * 4),
                    "d2h_bytes_per_step": int(img_pin.numel() * 4 + dep_pin.numel() * 4), "ms_per_step": ms_pose / args.steps,
                    "api": "ngp_render_poses_host (C ABI: host 4x4 pose in, rays generated in the render kernel's prologue, host image + depth out)"},
                "gpu_launches": int(launches),
                "roofline": roofline,
                "clocks": clocks.summary()}
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            n = args.cpu_rays or 100000
            v, dt, rays, _ = cpu_port_mrays(wl, args.variant, n)
            line["cpu_baseline"] = {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "port", "seconds": dt,
                                    "sample": f"{rays} rays (strided sub-sample of the same frame), oracle/ngp_oracle.c, OpenMP"}
            try:
                line["cpu_reference_run"] = cpu_run_sampler_mrays()
            except Exception as e:
                line["cpu_reference_run"] = {"unavailable": f"{type(e).__name__}: {e}"}
        if world == 1 and not args.no_ref_gpu:
            line["reference_gpu"] = reference_gpu(wl, w, bitfield, o_dev, d_dev, flush, dev)
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    return 0


def reference_gpu(wl, w, bitfield, o_dev, d_dev, flush, dev):
    """The reference's own extensions rebuilt for sm_100a (oracle/_ref), driven by its run_cuda loop, same frame, same box."""
    import torch
    try:
        from tests import refext
        if not refext.available("_ffmlp"):
            return {"unavailable": "oracle/_ref/*.so not built"}
        ref = refext.RefModel(w, bitfield, device=dev)
        r = ref.run_cuda(o_dev, d_dev)            # counted once (iterations, samples); the timed loop below is uninstrumented
        times = []
        for _ in range(3):
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ref.run_cuda(o_dev, d_dev, count=False)   # exactly the reference's loop: no per-ray counting, no extra host sync
            torch.cuda.synchronize()
            times.append(time.perf_counter() - t0)
        t = min(times)
        return {"value": o_dev.shape[0] / t / 1e6, "unit": "Mrays/s", "ms_per_step": t * 1e3, "iterations": r["iterations"],
                "samples": r["samples"], "what": "reference _raymarching/_gridencoder/_shencoder/_ffmlp rebuilt for sm_100a + run_cuda loop"}
    except Exception as e:  # the oracle is optional evidence, never the product
        return {"unavailable": f"{type(e).__name__}: {e}"}


if __name__ == "__main__":
    sys.exit(main())
