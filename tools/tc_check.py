#!/usr/bin/env python
"""Compares ngp_network_forward_tc (tcgen05 MLPs) with ngp_network_forward (mma.sync MLPs) on random points, and times both."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerfsafetyvalidation_b200 import _cabi, synthetic as syn  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    bitfield, grid = syn.stonehenge_bitfield()
    res = {}
    for variant in "AB":
        w = syn.ngp_weights(0, variant=variant)
        model = syn.build_model(w, bitfield, grid, device=dev)
        assert model._configure_native(model._native)
        h = model._native.get()
        lib = _cabi.load()
        for M in (1, 127, 128, 1000, 1 << 20):
            g = torch.Generator(device="cpu").manual_seed(M)
            x = (torch.rand(M, 3, generator=g) * 2 - 1).to(dev)
            d = torch.nn.functional.normalize(torch.randn(M, 3, generator=g), dim=-1).to(dev)
            out = []
            for name in ("ngp_network_forward", "ngp_network_forward_tc"):
                sg = torch.empty(M, device=dev)
                rgb = torch.empty(M, 3, device=dev)
                fn = getattr(lib, name)
                _cabi.check(fn(h, _cabi.ptr(x), _cabi.ptr(d), M, 1, _cabi.ptr(sg), _cabi.ptr(rgb), _cabi.stream()), name)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(5):
                    _cabi.check(fn(h, _cabi.ptr(x), _cabi.ptr(d), M, 1, _cabi.ptr(sg), _cabi.ptr(rgb), _cabi.stream()), name)
                e1.record()
                torch.cuda.synchronize()
                out.append((sg, rgb, e0.elapsed_time(e1) / 5))
            (s0, c0, t0), (s1, c1, t1) = out
            rel = ((s0 - s1).abs() / s0.abs().clamp_min(1e-6)).max().item()
            res[f"{variant}/{M}"] = {"sigma_rel": rel, "rgb_abs": (c0 - c1).abs().max().item(), "ms_mma_sync": t0, "ms_tcgen05": t1,
                                     "sigma_mean": s0.mean().item()}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
