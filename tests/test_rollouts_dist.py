"""CPU tests of the multi-GPU host logic: rollout sharding, the batched restatement of the drone
dynamics / pose conversion / get_rays, and the CEM elite-statistics update over a world_size-2 `gloo` group, which must give
every rank exactly what a single process computes with the reference's formulas (CrossEntropyMethod.py:211-251)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerfsafetyvalidation_b200 import rollouts as ro

STEPS = 20


def test_sharding_partitions_the_rollouts():
    for n, world in ((4096, 8), (2048, 4), (10, 3), (1, 2)):
        parts = [ro.shard_ids(n, r, world) for r in range(world)]
        allids = np.sort(np.concatenate(parts))
        assert np.array_equal(allids, np.arange(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_noise_is_a_function_of_the_rollout_id_only():
    a = ro.sample_noise(np.array([5, 9, 2]), STEPS)
    b = ro.sample_noise(np.array([2, 5]), STEPS)
    assert np.array_equal(a[0], b[1]) and np.array_equal(a[2], b[0])
    assert abs(a.std() / np.mean(ro.ENV["noise_std"]) - 1) < 0.2


def reference_drone_dynamics(state, action, dt, mass=1.0, g=10.0):
    """nav/agent_helpers.py:102-148 for ONE state, in torch like the reference (I = identity, envConfig.json:24)."""
    state, action = torch.as_tensor(state), torch.as_tensor(action)
    pos, v, R_flat, omega = state[0:3], state[3:6], state[6:9], state[9:]
    R = torch.as_tensor(ro.vec_to_rot_matrix(R_flat.numpy()))
    sum_action = torch.zeros(3, dtype=state.dtype)
    sum_action[2] = action[0]
    dv = (torch.tensor([0., 0., -mass * g], dtype=state.dtype) + R @ sum_action) / mass
    domega = action[1:] - torch.linalg.cross(omega, omega)
    angle = omega * dt
    theta = torch.norm(angle, p=2)
    exp_i = torch.eye(3, dtype=state.dtype)
    if theta != 0:
        K = torch.as_tensor(ro.skew((angle / theta).numpy()))
        exp_i = exp_i + torch.sin(theta) * K + (1 - torch.cos(theta)) * (K @ K)
    nxt = torch.zeros(12, dtype=state.dtype)
    nxt[0:3] = pos + v * dt
    nxt[3:6] = v + dv * dt
    nxt[6:9] = torch.as_tensor(ro.rot_matrix_to_vec((R @ exp_i).numpy()))
    nxt[9:] = omega + domega * dt
    return nxt.numpy()


def test_batched_dynamics_match_the_per_state_form():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((16, 12)) * 0.3
    x[0, 9:] = 0.0                                  # theta == 0 branch
    act = np.concatenate([np.full((16, 1), 10.0), rng.standard_normal((16, 3)) * 0.1], 1)
    got = ro.drone_dynamics(x, act, 0.1)
    for b in range(16):
        assert np.allclose(got[b], reference_drone_dynamics(x[b], act[b], 0.1), atol=1e-12)
    # rotation vector <-> matrix round trip
    v = rng.standard_normal((64, 3)) * 0.5
    assert np.allclose(ro.rot_matrix_to_vec(ro.vec_to_rot_matrix(v)), v, atol=1e-9)


def test_nominal_rollout_follows_the_straight_line():
    st = ro.rollout_states(np.zeros((1, STEPS, 12)))
    assert np.allclose(st[0, -1, 0:3], ro.ENV["end_pos"], atol=1e-9)      # hover thrust cancels gravity, v is constant
    poses = ro.states_to_ngp_poses(st[0])
    assert poses.shape == (STEPS, 4, 4) and np.allclose(poses[:, 3], [0, 0, 0, 1])
    R = poses[:, :3, :3].astype(np.float64)
    assert np.allclose(R @ np.swapaxes(R, -1, -2), np.eye(3), atol=1e-6) and np.allclose(np.linalg.det(R), 1, atol=1e-6)
    # NGP frame = (y, z, x) of the world frame (nav/math_utils.py:21-33)
    assert np.allclose(poses[-1, :3, 3], np.array(ro.ENV["end_pos"])[[1, 2, 0]], atol=1e-6)


def test_get_rays_matches_the_single_pose_generator():
    from nerfsafetyvalidation_b200 import synthetic as syn
    pose = syn.orbit_pose(80.0)
    o, d = ro.get_rays(pose[None], 30.0, 30.0, 6.0, 4.0, 8, 12)
    o1, d1 = syn.pinhole_rays(pose, 8, 12, focal=30.0)
    assert o.shape == (1, 96, 3) and np.allclose(o[0], o1) and np.allclose(d[0], d1, atol=1e-6)


def cem_reference(risks, noise, m_elite, pm, ps, qm, qs):
    """CrossEntropyMethod.py:215-251 with torch.distributions, single process."""
    from torch.distributions import MultivariateNormal as MVN
    idx = np.argsort(risks.numpy(), kind="stable")[:m_elite]
    es = noise[idx]
    means, variances = [], []
    for i in range(noise.shape[1]):
        p = MVN(torch.tensor(pm[i]), torch.diag(torch.tensor(ps[i]) ** 2))
        q = MVN(torch.tensor(qm[i]), torch.diag(torch.tensor(qs[i]) ** 2))
        lw = p.log_prob(es[:, i]) - q.log_prob(es[:, i])
        w = torch.exp(lw - torch.logsumexp(lw, 0))
        if torch.any(w <= 0):                      # CrossEntropyMethod.py:233-236
            w = torch.clamp(w, min=1e-8)
        means.append(w @ es[:, i])
        variances.append(torch.clamp(torch.cov(es[:, i].T, aweights=w).diag(), 0, 0.1))
    return torch.stack(means), torch.stack(variances), idx


def cem_inputs(n):
    noise = torch.from_numpy(ro.sample_noise(np.arange(n), STEPS, std=np.array(ro.ENV["noise_std"]) * 2.0))
    risks = torch.rand(n, generator=torch.Generator().manual_seed(1))
    pm = np.zeros((STEPS, 12))
    ps = np.broadcast_to(np.array(ro.ENV["noise_std"]), (STEPS, 12)).copy()
    qm = pm + 0.002
    qs = ps * 2.0
    return noise, risks, pm, ps, qm, qs


def test_cem_update_single_process():
    n, m = 96, 10
    noise, risks, pm, ps, qm, qs = cem_inputs(n)
    mean, var, elite = ro.cem_update(risks, noise, np.arange(n), n, m, pm, ps, qm, qs)
    rm, rv, ridx = cem_reference(risks, noise, m, pm, ps, qm, qs)
    assert sorted(elite) == sorted(ridx.tolist())
    assert torch.allclose(mean, rm, atol=1e-12) and torch.allclose(var, rv, atol=1e-12)
    assert (var >= 0).all() and (var <= 0.1).all()


def test_cem_update_with_one_dominant_elite():
    """A proposal far narrower than the nominal model makes the importance weights of all elites but a few underflow to exactly 0:
    the reference then clamps ALL weights to >= 1e-8 (CrossEntropyMethod.py:233-236).  The update must stay finite and equal to it."""
    n, m = 64, 8
    noise, risks, pm, ps, qm, qs = cem_inputs(n)
    qs = ps * 0.01                                  # log q ~ -(x / (0.01 std))^2 / 2: differences of thousands between elites
    rm, rv, ridx = cem_reference(risks, noise, m, pm, ps, qm, qs)
    lw_span = []
    mean, var, elite = ro.cem_update(risks, noise, np.arange(n), n, m, pm, ps, qm, qs)
    assert torch.isfinite(mean).all() and torch.isfinite(var).all() and (var >= 0).all() and (var <= 0.1).all()
    assert sorted(elite) == sorted(ridx.tolist())
    assert torch.allclose(mean, rm, atol=1e-12)
    assert torch.allclose(var, torch.nan_to_num(rv, nan=0.0, posinf=0.1, neginf=0.0), atol=1e-12)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _cem_worker(rank, world, port, n, m, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        noise, risks, pm, ps, qm, qs = cem_inputs(n)
        ids = ro.shard_ids(n, rank, world)
        mean, var, elite = ro.cem_update(risks[ids], noise[ids], ids, n, m, pm, ps, qm, qs, rank, world)
        torch.save(dict(mean=mean, var=var, elite=elite), os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,m", [(96, 10), (33, 7)])
def test_cem_update_world_size_2_gloo(tmp_path, n, m):
    world = 2
    mp.spawn(_cem_worker, args=(world, _free_port(), n, m, str(tmp_path)), nprocs=world, join=True)
    noise, risks, pm, ps, qm, qs = cem_inputs(n)
    rm, rv, ridx = cem_reference(risks, noise, m, pm, ps, qm, qs)
    outs = [torch.load(tmp_path / f"rank{r}.pt") for r in range(world)]
    for o in outs:
        assert sorted(o["elite"]) == sorted(ridx.tolist())
        assert torch.allclose(o["mean"], rm, atol=1e-12) and torch.allclose(o["var"], rv, atol=1e-12)
    assert torch.equal(outs[0]["mean"], outs[1]["mean"]) and torch.equal(outs[0]["var"], outs[1]["var"])
