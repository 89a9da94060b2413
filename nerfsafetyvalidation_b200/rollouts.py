"""Render-only disturbance rollouts and the CEM elite statistics, sharded over GPUs.

What validate.py's stress tests do around the hot path, reduced to what feeds and consumes NeRFRenderer.render:
  * poses      nominal straight-line trajectory start_pos -> end_pos (envConfig.json:32-33, planner_cfg), every step disturbed
               by a 12-D noise vector pushed through the drone dynamics (nav/agent_helpers.py:102-148) under hover thrust, then
               converted to the NGP camera frame exactly as Estimator.render_from_pose does (nav/estimator_helpers.py:227-235,
               nav/math_utils.py:21-33) and turned into rays like nerf/utils.py:get_rays (pixel centres + 0.5)
  * sharding   rollouts are independent (MonteCarlo.py:38, CrossEntropyMethod.py:79 are plain loops with per-rollout seeds):
               rollout i belongs to rank i % world, weights are replicated, there is NO data-path collective
  * CEM        per iteration: all-gather of the per-rollout risks (n_rollouts floats), every rank derives the same elite set,
               then ONE all-reduce of the importance-weighted moments of its own elites gives every rank the same
               mean / diagonal covariance update as CrossEntropyMethod.py:211-251 (clamp of the diagonal to [0, 0.1] included)

Host-side logic only (numpy + torch.distributed for the two small collectives; `gloo` on CPU in the tests, `nccl` on GPUs).
The Blender image fetch, SIFT estimator, planner and UQ of the real simulator step are out of scope (graphics / CPU control code).
"""
import contextlib
import math

import numpy as np

ENV = dict(  # envConfig.json (planner_cfg, agent_cfg, mpc_cfg)
    start_pos=(-0.75, -0.235, 0.25), end_pos=(0.2, -0.74, 0.3), T_final=2.0, mass=1.0, g=10.0,
    noise_mean=(0.0,) * 12, noise_std=(2e-2, 2e-2, 2e-2, 1e-2, 1e-2, 1e-2, 2e-2, 2e-2, 2e-2, 1e-2, 1e-2, 1e-2))


# ------------------------------------------------------------------ sharding
def shard_ids(n, rank, world):
    """rollout i -> rank i % world; returns this rank's ids in increasing order."""
    return np.arange(rank, n, world, dtype=np.int64)


# ------------------------------------------------------------------ rotations (nav/math_utils.py:152-190)
def skew(v):
    S = np.zeros(v.shape[:-1] + (3, 3), v.dtype)
    S[..., 0, 1], S[..., 0, 2] = -v[..., 2], v[..., 1]
    S[..., 1, 0], S[..., 1, 2] = v[..., 2], -v[..., 0]
    S[..., 2, 0], S[..., 2, 1] = -v[..., 1], v[..., 0]
    return S


def vec_to_rot_matrix(rot_vec):
    """Rodrigues, nav/math_utils.py:152-167 (axis = v / (1e-10 + |v|))."""
    angle = np.linalg.norm(rot_vec, axis=-1, keepdims=True)
    S = skew(rot_vec / (1e-10 + angle))
    a = angle[..., None]
    return np.eye(3, dtype=rot_vec.dtype) + np.sin(a) * S + (1 - np.cos(a)) * (S @ S)


def rot_matrix_to_vec(R):
    """Inverse of the above for rotations away from pi (log map)."""
    cos = np.clip((np.trace(R, axis1=-2, axis2=-1) - 1) / 2, -1.0, 1.0)
    angle = np.arccos(cos)
    v = np.stack([R[..., 2, 1] - R[..., 1, 2], R[..., 0, 2] - R[..., 2, 0], R[..., 1, 0] - R[..., 0, 1]], -1)
    s = np.sin(angle)
    k = np.where(np.abs(s) > 1e-8, angle / (2 * np.where(np.abs(s) > 1e-8, s, 1.0)), 0.5)
    return v * k[..., None]


# ------------------------------------------------------------------ dynamics (nav/agent_helpers.py:102-148), batched over rollouts
def drone_dynamics(state, action, dt, mass=1.0, g=10.0, inertia=None):
    """state [B,12] = pos, vel (world), rotation vector, body rates; action [B,4] = total thrust, body torques."""
    inertia = np.eye(3) if inertia is None else np.asarray(inertia, np.float64)
    pos, v, rvec, omega = state[:, 0:3], state[:, 3:6], state[:, 6:9], state[:, 9:12]
    R = vec_to_rot_matrix(rvec)
    thrust = np.zeros_like(pos)
    thrust[:, 2] = action[:, 0]
    dv = (np.array([0.0, 0.0, -mass * g]) + np.einsum("bij,bj->bi", R, thrust)) / mass
    domega = (action[:, 1:] - np.cross(omega, omega @ inertia.T)) @ np.linalg.inv(inertia).T
    angle = omega * dt
    theta = np.linalg.norm(angle, axis=-1)
    K = skew(angle / np.where(theta > 0, theta, 1.0)[:, None])
    exp_i = np.eye(3) + np.sin(theta)[:, None, None] * K + (1 - np.cos(theta))[:, None, None] * (K @ K)
    exp_i = np.where((theta == 0)[:, None, None], np.eye(3), exp_i)
    nxt = np.empty_like(state)
    nxt[:, 0:3] = pos + v * dt
    nxt[:, 3:6] = v + dv * dt
    nxt[:, 6:9] = rot_matrix_to_vec(R @ exp_i)
    nxt[:, 9:12] = omega + domega * dt
    return nxt


def sample_noise(rollout_ids, steps, mean=None, std=None):
    """[B, steps, 12] disturbance draws; generator seed = rollout id (SeedableMultivariateNormal.py:19-22), so the draw of
    rollout i does not depend on which rank owns it.  mean / std: [12] or per-step [steps, 12] (CEM proposal)."""
    mean = np.broadcast_to(np.asarray(ENV["noise_mean"] if mean is None else mean, np.float64), (steps, 12))
    std = np.broadcast_to(np.asarray(ENV["noise_std"] if std is None else std, np.float64), (steps, 12))
    out = np.empty((len(rollout_ids), steps, 12), np.float64)
    for b, rid in enumerate(rollout_ids):
        out[b] = mean + std * np.random.default_rng(int(rid)).standard_normal((steps, 12))
    return out


def rollout_states(noise, T_final=None):
    """noise [B, steps, 12] -> states [B, steps, 12]: x_{k+1} = f(x_k, hover) + noise_k (agent_helpers.py:43-77)."""
    B, steps, _ = noise.shape
    T = ENV["T_final"] if T_final is None else T_final
    dt = T / steps
    start, end = np.asarray(ENV["start_pos"], np.float64), np.asarray(ENV["end_pos"], np.float64)
    x = np.zeros((B, 12))
    x[:, 0:3] = start
    x[:, 3:6] = (end - start) / T
    hover = np.zeros((B, 4))
    hover[:, 0] = ENV["mass"] * ENV["g"]
    out = np.empty((B, steps, 12))
    for k in range(steps):
        x = drone_dynamics(x, hover, dt, ENV["mass"], ENV["g"]) + noise[:, k]
        out[:, k] = x
    return out


# ------------------------------------------------------------------ poses and rays
_ROT_X_90 = np.array([[1.0, 0.0, 0.0], [0.0, math.cos(math.pi / 2), -math.sin(math.pi / 2)], [0.0, math.sin(math.pi / 2), math.cos(math.pi / 2)]])
_NEG_YZ = np.diag([1.0, -1.0, -1.0])
_FLIP_YZ = np.array([[0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [1.0, 0.0, 0.0]])


def states_to_ngp_poses(states):
    """states [...,12] -> cam2world [...,4,4] in the NGP frame: rot_x(pi/2) @ R, then nerf_matrix_to_ngp_torch
    (nav/estimator_helpers.py:227-235, nav/math_utils.py:21-33)."""
    R = vec_to_rot_matrix(states[..., 6:9])
    rot = _FLIP_YZ @ (_ROT_X_90 @ R) @ _NEG_YZ
    trans = states[..., 0:3] @ _FLIP_YZ.T
    pose = np.zeros(states.shape[:-1] + (4, 4), np.float32)
    pose[..., :3, :3] = rot
    pose[..., :3, 3] = trans
    pose[..., 3, 3] = 1.0
    return pose


def get_rays(poses, fx, fy, cx, cy, H, W):
    """nerf/utils.py:52-116 for N = -1 (all pixels): poses [B,4,4] -> rays_o, rays_d float32 [B, H*W, 3]."""
    poses = np.asarray(poses, np.float32)
    i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    i = i.reshape(-1) + np.float32(0.5)
    j = j.reshape(-1) + np.float32(0.5)
    d = np.stack([(i - np.float32(cx)) / np.float32(fx), (j - np.float32(cy)) / np.float32(fy), np.ones_like(i)], -1)
    d = d / np.linalg.norm(d, axis=-1, keepdims=True).astype(np.float32)
    rays_d = np.einsum("nk,bjk->bnj", d, poses[:, :3, :3]).astype(np.float32)
    rays_o = np.broadcast_to(poses[:, None, :3, 3], rays_d.shape).astype(np.float32)
    return np.ascontiguousarray(rays_o), np.ascontiguousarray(rays_d)


def rollout_rays(rollout_ids, steps, H, W, focal, mean=None, std=None):
    """All frames of the given rollouts: (rays_o, rays_d) [B*steps, H*W, 3], states [B, steps, 12], noise [B, steps, 12]."""
    noise = sample_noise(rollout_ids, steps, mean, std)
    states = rollout_states(noise)
    poses = states_to_ngp_poses(states.reshape(-1, 12))
    o, d = get_rays(poses, focal, focal, W / 2, H / 2, H, W)
    return o, d, states, noise


# ------------------------------------------------------------------ CEM elite statistics (CrossEntropyMethod.py:211-251)
def _all_gather_1d(x, group_world):
    """x: 1-D tensor with the SAME length on every rank -> concatenation over ranks (rank order)."""
    import torch
    import torch.distributed as dist
    if group_world == 1:
        return x
    out = [torch.empty_like(x) for _ in range(group_world)]
    dist.all_gather(out, x)
    return torch.cat(out)


def cem_update(risks_local, noise_local, ids_local, n_rollouts, m_elite, p_mean, p_std, q_mean, q_std, rank=0, world=1):
    """One CEM proposal update, identical on every rank.

    risks_local [B_local] (float32 tensor), noise_local [B_local, steps, 12] (float64 tensor), ids_local [B_local] rollout
    ids owned by this rank (shard_ids).  p_*: nominal disturbance model, q_*: current proposal, each [steps, 12].
    Returns (new_mean [steps,12], new_var [steps,12], elite_ids) following CrossEntropyMethod.py:215-251: elites = the m_elite
    lowest risks; per step w = softmax(log p - log q) over the elites, mean = sum w x, var = diag of torch.cov(x.T,
    aweights=w) clamped to [0, 0.1].
    Collectives: one all-gather of the padded per-rank risks, one all-reduce(max) + one all-reduce(sum) of the moments.
    """
    import torch
    import torch.distributed as dist
    dev = risks_local.device
    per_rank = -(-n_rollouts // world)
    pad_r = torch.full((per_rank,), float("inf"), dtype=torch.float32, device=dev)
    pad_r[: risks_local.numel()] = risks_local.float()
    pad_i = torch.full((per_rank,), -1, dtype=torch.int64, device=dev)
    pad_i[: risks_local.numel()] = torch.as_tensor(ids_local, device=dev)
    all_r = _all_gather_1d(pad_r, world)
    all_i = _all_gather_1d(pad_i, world)
    # the same total order on every rank: by risk, ties by rollout id (np.argsort in the reference is order-dependent on ties).
    # ONE device->host copy of the two small vectors, then a host lexsort (no per-element .item() syncs)
    r_host, i_host = all_r.cpu().numpy(), all_i.cpu().numpy()
    order = np.lexsort((i_host, r_host))
    elite_ids = [int(i) for i in i_host[order[:m_elite]] if i >= 0]
    mine = torch.from_numpy(np.isin(np.asarray(ids_local, np.int64), np.asarray(elite_ids, np.int64))).to(dev)
    x = noise_local.to(dev)[mine].double()                                   # [E_local, steps, 12]
    steps = x.shape[1] if x.ndim == 3 else p_mean.shape[0]
    pm, ps, qm, qs = (torch.as_tensor(np.array(a, dtype=np.float64), device=dev) for a in (p_mean, p_std, q_mean, q_std))

    def logprob(v, m, s):   # diagonal MultivariateNormal.log_prob
        return (-0.5 * ((v - m) / s) ** 2 - torch.log(s) - 0.5 * math.log(2 * math.pi)).sum(-1)

    lw = logprob(x, pm, ps) - logprob(x, qm, qs) if x.numel() else torch.empty(0, steps, dtype=torch.float64, device=dev)   # [E_local, steps]
    mx = lw.max(dim=0).values if lw.numel() else torch.full((steps,), -float("inf"), dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    # CrossEntropyMethod.py:225-236: w = exp(lw - logsumexp(lw)); if ANY weight underflowed to <= 0, w = clamp(w, min=1e-8).
    # With e = exp(lw - max) and S = sum e over all elites, w = e / S, and w <= 0 exactly where e == 0: the first all-reduce
    # carries S, the moments of e and the global count of underflowed weights.  Only if that count is non-zero (rare) a second
    # all-reduce redoes the moments with the clamped weights, as the reference would.
    e = torch.exp(lw - mx)                                                   # unnormalised, <= 1

    def moments(wt, extra=None):   # sum w, sum w^2, sum w x [12], sum w x^2 [12] (, extra)
        m = torch.zeros(steps, 27, dtype=torch.float64, device=dev)
        if wt.numel():
            m[:, 0] = wt.sum(0)
            m[:, 1] = (wt * wt).sum(0)
            m[:, 2:14] = torch.einsum("es,esk->sk", wt, x)
            m[:, 14:26] = torch.einsum("es,esk->sk", wt, x * x)
            if extra is not None:
                m[:, 26] = extra.sum(0)
        if world > 1:
            dist.all_reduce(m, op=dist.ReduceOp.SUM)
        return m

    mom = moments(e, (e <= 0).double())
    S = mom[:, 0:1]
    if bool((mom[:, 26] > 0).any()):                                         # identical on every rank (all-reduced)
        mom = moments(torch.clamp(e / S.t(), min=1e-8))
        sw, sw2, swx, swx2 = mom[:, 0:1], mom[:, 1:2], mom[:, 2:14], mom[:, 14:26]
    else:
        sw, sw2, swx, swx2 = S / S, mom[:, 1:2] / (S * S), mom[:, 2:14] / S, mom[:, 14:26] / S
    mean = swx                                                               # :238  mean = weights @ samples (weights as clamped, not renormalised)
    avg = swx / sw                                                           # torch.cov normalises its aweights itself
    # torch.cov(aweights=w, correction=1): sum w (x - avg)^2 / (sum w - sum w^2 / sum w)
    denom = sw - sw2 / sw
    var = (swx2 - sw * avg * avg) / torch.where(denom > 0, denom, torch.full_like(denom, float("nan")))
    var = torch.nan_to_num(var, nan=0.0, posinf=0.1, neginf=0.0)             # one elite carrying all the weight: torch.cov gives nan/inf; the proposal keeps a zero width there
    var = torch.clamp(var, 0.0, 0.1)                                         # :243-245
    return mean, var, elite_ids


# ------------------------------------------------------------------ GPU drivers (one process per GPU; weights replicated)
def get_rays_device(poses, fx, fy, cx, cy, H, W):
    """get_rays (nerf/utils.py:52-116) on poses' device with the ngp_get_rays kernel: [B,4,4] -> rays_o, rays_d [B, H*W, 3] fp32."""
    from .nerf.utils import get_rays as get_rays_cuda
    r = get_rays_cuda(poses, (fx, fy, cx, cy), H, W)
    return r["rays_o"], r["rays_d"]


def render_rollouts(model, ids, steps, H, W, focal, frames_per_call=64, mean=None, std=None, **render_kw):
    """Renders every frame of this rank's rollouts `ids` with model.render (frames batched as [F, H*W, 3] per call, so the
    batch shares one n_step schedule exactly like renderer.py:265-269 flattens it) and returns
    (risks [len(ids)] float32 on the device, noise [len(ids), steps, 12] float64 tensor, rays rendered)."""
    import torch
    dev = next(model.parameters()).device
    noise = sample_noise(ids, steps, mean, std)
    poses = torch.from_numpy(states_to_ngp_poses(rollout_states(noise).reshape(-1, 12))).to(dev)
    risks = []
    centre = (H // 2) * W + W // 2
    intr = (focal, focal, W / 2, H / 2)
    frozen = model.frozen_weights() if hasattr(model, "frozen_weights") else contextlib.nullcontext()
    with torch.no_grad(), frozen:
        for f0 in range(0, poses.shape[0], frames_per_call):
            # 64 B per frame in: the rays (get_rays, nerf/utils.py:52-116) are generated inside the render kernel
            out = model.render_poses(poses[f0:f0 + frames_per_call], intr, H, W, bg_color=1, perturb=False, **render_kw)
            F = out["depth"].shape[0]
            # per-frame clearance along the optical axis: normalised depth of what the centre ray hit, pushed to 1 where it
            # hit nothing (weights_sum ~ 0); missed rays have NaN depth (renderer.py:376)
            dc = torch.nan_to_num(out["depth"][:, centre], nan=1.0)
            wc = out["weights_sum"].view(F, H * W)[:, centre]
            risks.append((dc + (1.0 - wc)).clamp(0.0, 1.0))
    per_frame = torch.cat(risks).view(len(ids), steps)
    return per_frame.amin(dim=1).float(), torch.from_numpy(noise), poses.shape[0] * H * W


def monte_carlo(model, n_rollouts, steps, H, W, focal, rank=0, world=1, **kw):
    """cfg3: this rank's share of a Monte Carlo stress test.  No collective on the data path; the caller may all-gather the
    returned risks (one float per rollout) for reporting."""
    ids = shard_ids(n_rollouts, rank, world)
    risks, _, rays = render_rollouts(model, ids, steps, H, W, focal, **kw)
    return ids, risks, rays


def cross_entropy_method(model, n_iters, n_rollouts, steps, H, W, focal, m_elite=None, rank=0, world=1, **kw):
    """cfg4: CEM over the per-step disturbance proposal (validate.py:31-38 initialises q = p = N(noise_mean, diag(std^2))).
    Every iteration: render this rank's rollouts under q, then cem_update (two tiny collectives) -> the same q on all ranks."""
    m_elite = m_elite or max(1, n_rollouts // 10)
    p_mean = np.broadcast_to(np.asarray(ENV["noise_mean"], np.float64), (steps, 12)).copy()
    p_std = np.broadcast_to(np.asarray(ENV["noise_std"], np.float64), (steps, 12)).copy()
    q_mean, q_std = p_mean.copy(), p_std.copy()
    history, rays_total = [], 0
    for k in range(n_iters):
        # rollout ids are offset per iteration so every population draws fresh noise (seed = id)
        ids = shard_ids(n_rollouts, rank, world) + k * n_rollouts
        risks, noise, rays = render_rollouts(model, ids, steps, H, W, focal, mean=q_mean, std=q_std, **kw)
        mean, var, elite = cem_update(risks, noise, ids, n_rollouts, m_elite, p_mean, p_std, q_mean, q_std, rank, world)
        q_mean, q_std = mean.cpu().numpy(), np.sqrt(np.maximum(var.cpu().numpy(), 1e-12))
        rays_total += rays
        history.append(dict(iteration=k, mean_risk_local=float(risks.mean()), n_elite=len(elite)))
    return q_mean, q_std, history, rays_total
