"""Latency of the differentiable small-batch render on the estimator's shape: nav/estimator_helpers.py:
191-225 renders <= 1024 rays per iteration and calls loss.backward() on the pose, 100 iterations per filter step.

Prints one JSON line: forward / backward / forward+backward milliseconds of `model.render` with rays that require grad (CUDA events on
the current stream, median of --iters after --warmup), the same for `model.density` on the planner's shape (steps x 500 points), and --
when oracle/_ref holds the rebuilt reference extensions -- the reference's own training-mode pieces (march_rays_train -> its network
under autocast -> composite_rays_train, tests/test_gpu_diff.reference_parts_grads) on the same box for the same rays.

    python tools/diff_bench.py [--rays 1024] [--variant B] [--iters 50]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerfsafetyvalidation_b200 import _cabi, synthetic as syn   # noqa: E402


def timed(fn, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=1024)
    ap.add_argument("--points", type=int, default=20 * 500)
    ap.add_argument("--variant", default="B")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    cuda = torch.device("cuda:0")
    bitfield, grid = syn.stonehenge_bitfield()
    w = syn.ngp_weights(0, variant=args.variant)
    model = syn.build_model(w, bitfield, grid, device=cuda)
    o, d = syn.pinhole_rays(syn.orbit_pose(), 128, 128)
    pick = np.random.default_rng(0).permutation(o.shape[0])[: args.rays]
    o = torch.from_numpy(np.ascontiguousarray(o[pick], np.float32)).to(cuda)
    d = torch.from_numpy(np.ascontiguousarray(d[pick], np.float32)).to(cuda)
    target = torch.rand(args.rays, 3, device=cuda)
    kw = dict(staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024)
    state = {}

    def fwd():
        state["o"], state["d"] = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
        state["loss"] = torch.nn.functional.mse_loss(model.render(state["o"][None], state["d"][None], **kw)["image"][0], target)

    def bwd():
        state["loss"].backward()

    def both():
        fwd()
        bwd()

    with model.frozen_weights():
        launches0 = _cabi.launch_count()
        both()
        launches = _cabi.launch_count() - launches0
        with torch.no_grad():
            t_inf = timed(lambda: model.render(o[None], d[None], **kw), args.iters, args.warmup)
        t_fwd = timed(fwd, args.iters, args.warmup)
        t_both = timed(both, args.iters, args.warmup)
        x = (torch.rand(args.points, 3, device=cuda) * 2 - 1)

        def dens():
            xr = x.clone().requires_grad_(True)
            model.density(xr)["sigma"].sum().backward()
        t_dens = timed(dens, args.iters, args.warmup)
        spr = model.render(o[None], d[None], return_stats=True, **kw)["samples_per_ray"].float().mean().item()
    out = dict(workload=f"{args.rays} rays of the cfg2 orbit view, variant {args.variant}, max_steps 1024; density: {args.points} points",
               mean_samples_per_ray=spr, inference_render_ms=t_inf, diff_forward_ms=t_fwd, diff_forward_backward_ms=t_both,
               diff_backward_ms=t_both - t_fwd, density_forward_backward_ms=t_dens, kernel_launches_forward_backward=int(launches))
    try:
        from tests import refext
        from tests.test_gpu_diff import build_reference_model, reference_parts_grads
        if refext.reference_training_stack() is not None:
            G = torch.randn(args.rays, 3, device=cuda)
            g_ws = torch.zeros(args.rays, device=cuda)
            ref, rm = build_reference_model(w, model, cuda)
            out["reference_pieces_forward_backward_ms"] = timed(lambda: reference_parts_grads(ref, rm, o, d, G, g_ws), max(args.iters // 5, 3), 2)
    except Exception as e:   # the reference side is optional evidence
        out["reference_pieces_error"] = repr(e)[:200]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
