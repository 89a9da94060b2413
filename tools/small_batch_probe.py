"""Times the two renderers on small ray batches: the persistent cooperative kernel (ngp_render) and the warp-per-ray forward
(ngp_render_diff_forward), CUDA events, median of 30.  python tools/small_batch_probe.py"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerfsafetyvalidation_b200 import _cabi, synthetic as syn   # noqa: E402
from tools.diff_bench import timed   # noqa: E402

cuda = torch.device("cuda:0")
bitfield, grid = syn.stonehenge_bitfield()
out = {}
for var in "AB":
    model = syn.build_model(syn.ngp_weights(0, variant=var), bitfield, grid, device=cuda)
    o, d = syn.pinhole_rays(syn.orbit_pose(), 256, 256)
    lib = _cabi.load()
    with model.frozen_weights(), torch.no_grad():
        for n in (1024, 4096, 16384, 40000, 65536):
            pick = np.random.default_rng(0).permutation(o.shape[0])[:n]
            oo = torch.from_numpy(np.ascontiguousarray(o[pick], np.float32)).to(cuda)
            dd = torch.from_numpy(np.ascontiguousarray(d[pick], np.float32)).to(cuda)
            kw = dict(staged=True, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024)
            a = timed(lambda: model.render(oo[None], dd[None], **kw), 30, 5)
            img, dep, ws, fin = (torch.empty(n, 3, device=cuda), torch.empty(n, device=cuda), torch.empty(n, device=cuda),
                                 torch.empty(n, 4, device=cuda))
            h = model._native.get(cuda)

            def warp_per_ray():
                _cabi.check(lib.ngp_render_diff_forward(h, _cabi.ptr(oo), _cabi.ptr(dd), n, 0.0, 1024, 1.0, _cabi.ptr(img), _cabi.ptr(dep),
                                                        _cabi.ptr(ws), _cabi.ptr(fin), None, _cabi.stream()))
            b = timed(warp_per_ray, 30, 5)
            ref = model.render(oo[None], dd[None], **kw)["image"][0]
            out[f"{var}_{n}"] = dict(persistent_ms=a, warp_per_ray_ms=b, max_image_diff=float((ref - img).abs().max()))
print(json.dumps(out))
