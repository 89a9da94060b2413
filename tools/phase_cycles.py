#!/usr/bin/env python
"""Per-phase warp-cycle breakdown of ngp_render on the bench frame (diagnostics; needs a GPU).

    python tools/phase_cycles.py [--workload cfg2] [--variant A]
"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from nerfsafetyvalidation_b200 import _cabi, synthetic as syn  # noqa: E402

NAMES = ["other (tile set-up / end, prologue, epilogue)", "march phase (all 32 warps)", "barrier after march", "gather warps: sample load + gather + operand store",
         "epilogue warps: activations + compositing", "gather warps waiting (operand slot, accumulators)", "epilogue warps waiting (layer commits)",
         "barrier after evaluate"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--variant", default="A")
    args = ap.parse_args()
    wl = bench.WORKLOADS[args.workload]
    dev = torch.device("cuda:0")
    bitfield, grid = bench.scene_for(wl)
    w = syn.ngp_weights(0, log2_hashmap_size=wl["log2T"], variant=args.variant)
    model = syn.build_model(w, bitfield, grid, device=dev)
    o, d = bench.frame_rays(wl, 0, wl.get("frames", 1))
    o, d = torch.from_numpy(o).to(dev)[None], torch.from_numpy(d).to(dev)[None]
    lib = _cabi.load()
    with torch.no_grad():
        out = model.render(o, d, bg_color=1, perturb=False, return_stats=True)
        torch.cuda.synchronize()
        stats = out["stats"].cpu().tolist()
        h = model._native.get()
        _cabi.check(lib.ngp_ctx_enable_profile(h, 1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        model.render(o, d, bg_color=1, perturb=False)
        e1.record()
        torch.cuda.synchronize()
        buf = (C.c_int64 * 8)()
        _cabi.check(lib.ngp_ctx_phase_cycles(h, buf, 1))
    cyc = np.array(list(buf), dtype=np.float64)
    n_warps = torch.cuda.get_device_properties(dev).multi_processor_count * 32   # 16 gather + 16 epilogue warps per SM
    chunks = stats[1] / 32
    total = cyc.sum()   # slot 0 absorbs whatever the others did not count: the eight slots add up to every warp's lifetime
    res = {"ms_profiled_launch": e0.elapsed_time(e1), "iterations": stats[0], "samples": stats[1], "epochs": stats[2], "replays": stats[3],
           "phases_pct_of_warp_time": {n: round(100 * c / total, 2) for n, c in zip(NAMES, cyc)},
           "avg_cycles_per_warp": {n: round(c / n_warps) for n, c in zip(NAMES, cyc)},
           "warp_cycles_per_32_samples": {n: round(c / chunks) for n, c in zip(NAMES, cyc)}}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
