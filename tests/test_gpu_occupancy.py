"""GPU: occupancy-grid maintenance against the reference's OWN NeRFRenderer methods
(/root/reference/nerf/renderer.py:388-546, copied unmodified to oracle/_ref/pyref by oracle/build_ref.py and run on the
reference's extensions rebuilt for sm_100a), both driven from the same torch CUDA generator seed:

  update_extra_state, full branch     density grid and bitfield after one and after three calls
  update_extra_state, partial branch  (iter_density >= 16) the same random cells are queried; torch's scatter picks an arbitrary
                                      winner among duplicate indices upstream, so cells that were drawn once are compared exactly
                                      and the rest statistically
  mark_untrained_grid                 the -1 mask of cells no camera covers
Density values differ between the two only by the MLP accumulation (fp32 here, fp16 in the reference's FFMLP): <= 2e-2 relative,
which moves a cell across the bitfield threshold only when its density sits within that distance of it.
"""
import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import synthetic as syn
from tests import refext

pytestmark = pytest.mark.gpu
RefRenderer = refext.reference_renderer_class() if refext.available("_ffmlp") else None
needs_ref = pytest.mark.skipif(RefRenderer is None, reason="oracle/_ref/pyref + oracle/_ref/*.so not built (python oracle/build_ref.py)")


def make_pair(cuda, scene, variant="B"):
    bitfield, grid = scene
    w = syn.ngp_weights(0, variant=variant)
    ours = syn.build_model(w, bitfield, grid, device=cuda)
    net = refext.RefModel(w, bitfield, device=cuda)

    class Ref(RefRenderer):
        def density(self, x):
            return net.density(x)

    ref = Ref(bound=w["bound"], cuda_ray=True, density_scale=w["density_scale"], min_near=ours.min_near, density_thresh=ours.density_thresh).to(cuda)
    ref.density_grid.copy_(ours.density_grid)
    ref.density_bitfield.copy_(ours.density_bitfield)
    ref.eval()
    return ours, ref


def grids_agree(ours, ref, what):
    go, gr = ours.density_grid, ref.density_grid
    assert torch.equal(go < 0, gr < 0), f"{what}: untrained (-1) cells"
    live = gr >= 0
    rel = ((go - gr).abs() / gr.abs().clamp_min(1e-6))[live]
    assert float(rel.max()) <= 2e-2, f"{what}: density grid, max relative difference {float(rel.max()):.3e}"
    thresh = min(ref.mean_density, ref.density_thresh)
    assert abs(ours.mean_density - ref.mean_density) <= 2e-2 * abs(ref.mean_density) + 1e-9
    diff = ours.density_bitfield != ref.density_bitfield
    if bool(diff.any()):   # every differing bit belongs to a cell whose density is within the MLP tolerance of the threshold
        bits = torch.nonzero(diff).flatten()
        cells = (bits[:, None] * 8 + torch.arange(8, device=bits.device)[None, :]).flatten()
        flipped = ((go.flatten()[cells] > min(ours.mean_density, ours.density_thresh)) != (gr.flatten()[cells] > thresh))
        near = (gr.flatten()[cells] - thresh).abs() <= 4e-2 * abs(thresh) + 1e-9
        assert bool((~flipped | near).all()), f"{what}: a bitfield bit differs away from the threshold"
        assert int(flipped.sum()) <= max(8, go.numel() // 2000), f"{what}: {int(flipped.sum())} cells flipped"


@needs_ref
def test_update_extra_state_full_branch_reproduces_the_reference_from_the_same_seed(cuda, scene):
    ours, ref = make_pair(cuda, scene)
    for call in range(3):
        torch.manual_seed(1234 + call)
        ref.update_extra_state()
        torch.manual_seed(1234 + call)
        ours.update_extra_state()
        assert ours.iter_density == ref.iter_density == call + 1
        grids_agree(ours, ref, f"full update, call {call}")
    # a smaller block size changes the order in which the reference consumes the generator; ours follows it
    torch.manual_seed(77)
    ref.update_extra_state(S=64)
    torch.manual_seed(77)
    ours.update_extra_state(S=64)
    grids_agree(ours, ref, "full update, S=64")


@needs_ref
def test_update_extra_state_partial_branch(cuda, scene):
    ours, ref = make_pair(cuda, scene)
    ours.iter_density = ref.iter_density = 16
    ours.local_step = ref.local_step = 4
    ours.step_counter[:4, 0] = ref.step_counter[:4, 0] = torch.tensor([10, 20, 30, 44], dtype=torch.int32, device=cuda)
    before = ref.density_grid.clone()
    torch.manual_seed(99)
    ref.update_extra_state()
    torch.manual_seed(99)
    ours.update_extra_state()
    assert ours.mean_count == ref.mean_count == 26 and ours.local_step == ref.local_step == 0
    go, gr = ours.density_grid, ref.density_grid
    touched = gr != before
    assert 0.2 < float(touched.float().mean()) < 0.9
    # the same cells were queried (same generator stream): a cell the reference left alone is left alone by us as well
    assert float(((go != before) == touched).float().mean()) >= 0.999
    rel = ((go - gr).abs() / gr.abs().clamp_min(1e-6))
    # duplicates: torch.Tensor.__setitem__ with repeated indices keeps an arbitrary one upstream; everything else within the MLP tolerance
    assert float((rel <= 2e-2).float().mean()) >= 0.97
    assert float((ours.density_bitfield == ref.density_bitfield).float().mean()) >= 0.995


@needs_ref
def test_mark_untrained_grid(cuda, scene):
    from nerfsafetyvalidation_b200.nerf.network_ff import NeRFNetwork
    ours = NeRFNetwork(bound=2, cuda_ray=True).to(cuda).eval()      # two cascades; the network itself is not involved
    ref = RefRenderer(bound=2, cuda_ray=True).to(cuda).eval()
    poses = np.stack([syn.orbit_pose(azimuth_deg=a, radius=1.6) for a in (0.0, 40.0, 95.0)]).astype(np.float32)
    intrinsic = (300.0, 300.0, 100.0, 100.0)
    ours.density_grid.fill_(0.5)
    ref.density_grid.fill_(0.5)
    ref.mark_untrained_grid(poses, intrinsic)
    ours.mark_untrained_grid(poses, intrinsic)
    a, b = ours.density_grid < 0, ref.density_grid < 0
    assert 0.02 < float(b.float().mean()) < 0.98, "the test cameras must see part of the grid, not all or nothing"
    # a cell on the edge of a frustum may fall either way (3-term dot products in a different order than cuBLAS' bmm)
    assert float((a != b).float().mean()) <= 1e-4
    assert torch.equal(ours.density_grid[~a], ref.density_grid[~a])
