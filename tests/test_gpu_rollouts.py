"""GPU tests of the rollout drivers (cfg3 / cfg4 shape, tiny sizes): frames batched per render call, risk from the rendered
depth, CEM proposal update.  World size 1 here; the collectives are covered on CPU by tests/test_rollouts_dist.py."""
import numpy as np
import pytest
import torch

from nerfsafetyvalidation_b200 import rollouts as ro, synthetic as syn

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model(cuda, scene):
    bitfield, grid = scene
    return syn.build_model(syn.ngp_weights(0, variant="B"), bitfield, grid, device=cuda)


def test_get_rays_device_matches_host(cuda):
    st = ro.rollout_states(ro.sample_noise(np.arange(3), 5))
    poses = ro.states_to_ngp_poses(st.reshape(-1, 12))
    o, d = ro.get_rays(poses, 40.0, 40.0, 16.0, 12.0, 24, 32)
    od, dd = ro.get_rays_device(torch.from_numpy(poses).to(cuda), 40.0, 40.0, 16.0, 12.0, 24, 32)
    assert torch.allclose(od.cpu(), torch.from_numpy(o)) and torch.allclose(dd.cpu(), torch.from_numpy(d), atol=1e-6)


def test_monte_carlo_rollouts_render_and_batching(model, cuda):
    ids, risks, rays = ro.monte_carlo(model, 6, 4, 32, 32, 44.4, frames_per_call=8)
    assert rays == 6 * 4 * 32 * 32 and risks.shape == (6,) and torch.isfinite(risks).all() and (risks >= 0).all() and (risks <= 1).all()
    # the same rollouts rendered one frame per call: the schedule differs (renderer.py:357 is global over the batch), so the
    # depths agree to tolerance, not bitwise
    _, risks1, _ = ro.monte_carlo(model, 6, 4, 32, 32, 44.4, frames_per_call=1)
    assert (risks - risks1).abs().max().item() <= 1e-3
    # sharded over two ranks = the same rollouts
    a = ro.monte_carlo(model, 6, 4, 32, 32, 44.4, rank=0, world=2, frames_per_call=8)
    b = ro.monte_carlo(model, 6, 4, 32, 32, 44.4, rank=1, world=2, frames_per_call=8)
    merged = torch.empty(6, device=cuda)
    merged[torch.as_tensor(a[0])] = a[1]
    merged[torch.as_tensor(b[0])] = b[1]
    assert (merged - risks).abs().max().item() <= 1e-3


def test_cross_entropy_method_updates_the_proposal(model):
    q_mean, q_std, hist, rays = ro.cross_entropy_method(model, 2, 12, 3, 24, 24, 33.3, m_elite=4, frames_per_call=12)
    assert q_mean.shape == (3, 12) and q_std.shape == (3, 12) and len(hist) == 2 and rays == 2 * 12 * 3 * 24 * 24
    assert np.isfinite(q_mean).all() and (q_std >= 0).all() and (q_std ** 2 <= 0.1 + 1e-12).all()
    assert np.abs(q_mean).max() > 0            # the proposal moved away from the nominal zero mean
    again = ro.cross_entropy_method(model, 2, 12, 3, 24, 24, 33.3, m_elite=4, frames_per_call=12)
    assert np.array_equal(again[0], q_mean)    # deterministic: seeds are rollout ids
