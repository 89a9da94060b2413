#!/usr/bin/env python
"""How much of the 1e-3 image tolerance does the fused renderer use against the reference's rebuilt extensions?  (needs a GPU and
oracle/_ref/*.so; prints max |diff| of image / depth / weights_sum and the number of rays whose sample count differs)"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerfsafetyvalidation_b200 import synthetic as syn  # noqa: E402
from tests import refext  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    bitfield, grid = syn.stonehenge_bitfield()
    res = {}
    for variant in "AB":
        w = syn.ngp_weights(0, variant=variant)
        o, d = syn.pinhole_rays(syn.orbit_pose(), 400, 400)
        o, d = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
        ref = refext.RefModel(w, bitfield, device=dev).run_cuda(o, d)
        with torch.no_grad():
            out = syn.build_model(w, bitfield, grid, device=dev).render(o[None], d[None], bg_color=1, perturb=False, return_stats=True)
        img, dep, ws = out["image"].reshape(-1, 3), out["depth"].reshape(-1), out["weights_sum"].reshape(-1)
        hit = ~torch.isnan(dep)
        res[variant] = {"image": (img - ref["image"]).abs().max().item(), "depth": (dep - ref["depth"])[hit].abs().max().item(),
                        "weights_sum": (ws - ref["weights_sum"]).abs().max().item(),
                        "image_mean_abs": (img - ref["image"]).abs().mean().item(),
                        "rays_with_other_sample_count": int((out["samples_per_ray"] != ref["samples_per_ray"]).sum()), "rays": int(img.shape[0])}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
