#!/usr/bin/env python
"""Monte Carlo / Cross Entropy Method stress-test shapes (BASELINE.json configs[2], [3]) on 1..8 GPUs.

    python tools/stress_test.py --mode mc --rollouts 4096 --steps 20 --res 200
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/stress_test.py --mode cem --iters 8 --rollouts 2048

One process per GPU; rank 0 builds the weights and broadcasts them once (NCCL); rollouts shard rank = id % world with no
data-path collective; CEM adds one all-gather + two all-reduces of a few KB per iteration (rollouts.cem_update).
Prints one JSON line with Mrays/s (max-over-ranks device time)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from nerfsafetyvalidation_b200 import rollouts as ro, synthetic as syn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="mc", choices=["mc", "cem"])
    ap.add_argument("--rollouts", type=int, default=256)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--res", type=int, default=200)
    ap.add_argument("--frames-per-call", type=int, default=64)
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    bitfield, grid = syn.stonehenge_bitfield()
    model = syn.build_model(syn.ngp_weights(0 if rank == 0 else 1, variant="B"), bitfield, grid, device=dev)
    if world > 1:   # replicated weights: ONE broadcast per tensor at start-up
        for t in (model.encoder.embeddings.data, model.sigma_net.weights.data, model.color_net.weights.data, model.density_bitfield):
            dist.broadcast(t, src=0)
        model.invalidate_caches()
    focal = 277.78 * args.res / 200
    kw = dict(rank=rank, world=world, frames_per_call=args.frames_per_call)
    ro.monte_carlo(model, world * 4, 2, args.res, args.res, focal, **kw)   # warm-up
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if args.mode == "mc":
        _, risks, rays = ro.monte_carlo(model, args.rollouts, args.steps, args.res, args.res, focal, **kw)
        extra = {"mean_risk_rank0": float(risks.mean())}
    else:
        q_mean, q_std, hist, rays = ro.cross_entropy_method(model, args.iters, args.rollouts, args.steps, args.res, args.res, focal, **kw)
        extra = {"final_q_mean_abs_max": float(abs(q_mean).max()), "history_rank0": hist[-1]}
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1), float(rays)], dtype=torch.float64, device=dev)
    if world > 1:
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms, rays_all = float(tm[0]), float(t[1])
    else:
        ms, rays_all = float(t[0]), float(t[1])
    if rank == 0:
        print(json.dumps({"mode": args.mode, "n_gpus": world, "rollouts": args.rollouts, "steps": args.steps, "res": args.res,
                          "iters": args.iters if args.mode == "cem" else None, "rays": rays_all, "ms": ms,
                          "mrays_per_s": rays_all / ms / 1e3, "includes": "host pose generation, device ray generation, render, risk", **extra}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
