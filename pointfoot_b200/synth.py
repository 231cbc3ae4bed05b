"""Seeded synthetic simulation state (stands in for what PhysX would hand the env).

Isaac Gym is a closed binary that cannot be installed here, so ``gym.simulate`` is replaced on BOTH the
reference/oracle side and the B200 side by the same deterministic tensors.
Layouts are the ones PhysX owns (AoS): root_states[N,13], dof_state[N*D,2], contact_forces[N*B,3],
rigid_body_states[N*B,13]  (legged_gym/envs/pointfoot/point_foot.py:713-724).
"""
import numpy as np
import torch

NUM_DOF = 6
NUM_BODIES = 9
FEET = (4, 8)
# soft range the synthetic dof positions are drawn from (URDF limits, PF_P441A.urdf:148,191,234,390,433,476)
DOF_LOWER = np.array([-0.38397, -1.012291, -0.872665, -1.39626, -1.518436, -1.396263], np.float32)
DOF_UPPER = np.array([1.39626, 1.518436, 1.396263, 0.38397, 1.012291, 0.872665], np.float32)

TERRAIN_ROWS, TERRAIN_COLS = 10, 20          # pointfoot_rough_config.py:32-33
TERRAIN_LEN_PX = 80                          # 8 m / 0.1 m
BORDER_PX = 250                              # 25 m / 0.1 m


def make_heightfield(seed=1):
    """int16[1300,2100]: stairs (+-46 units, 3-cell treads), slopes and +-10 noise; 250-cell zero border.
    Returns (height_samples int16, terrain_origins f32[10,20,3])."""
    rng = np.random.default_rng(seed + 7919)
    rows = TERRAIN_ROWS * TERRAIN_LEN_PX + 2 * BORDER_PX
    cols = TERRAIN_COLS * TERRAIN_LEN_PX + 2 * BORDER_PX
    hf = np.zeros((rows, cols), np.int16)
    origins = np.zeros((TERRAIN_ROWS, TERRAIN_COLS, 3), np.float32)
    ii, jj = np.meshgrid(np.arange(TERRAIN_LEN_PX), np.arange(TERRAIN_LEN_PX), indexing="ij")
    ring = np.maximum(np.abs(ii - 39.5), np.abs(jj - 39.5))       # chebyshev distance from tile centre
    for i in range(TERRAIN_ROWS):
        for j in range(TERRAIN_COLS):
            kind = j % 4
            diff = (i + 1) / TERRAIN_ROWS
            if kind == 0:      # pyramid stairs up
                tile = (np.floor((39.5 - ring) / 3.0) * (46 * diff)).astype(np.int32)
            elif kind == 1:    # pyramid stairs down
                tile = -(np.floor((39.5 - ring) / 3.0) * (46 * diff)).astype(np.int32)
            elif kind == 2:    # slope + noise
                tile = ((39.5 - ring) * 8 * diff).astype(np.int32) + rng.integers(-10, 11, ring.shape)
            else:              # rough flat
                tile = rng.integers(-10, 11, ring.shape) * (1 + i // 3)
            r0 = BORDER_PX + i * TERRAIN_LEN_PX
            c0 = BORDER_PX + j * TERRAIN_LEN_PX
            hf[r0:r0 + TERRAIN_LEN_PX, c0:c0 + TERRAIN_LEN_PX] = tile.astype(np.int16)
            origins[i, j] = ((i + 0.5) * 8.0, (j + 0.5) * 8.0,
                             float(tile[38:42, 38:42].max()) * 0.005)
    return torch.from_numpy(hf), torch.from_numpy(origins)


def make_state(num_envs, task="pointfoot_flat", seed=1, env_id_offset=0, adversarial=False,
               num_reward_terms=19, terrain=None):
    """Dict of CPU tensors.  Env ``i`` of a shard with ``env_id_offset=o`` equals env ``o+i`` of the unsharded
    state only for the fields that are pure functions of the global id (terrain_types); everything else is
    drawn per shard from ``seed + env_id_offset``."""
    N, D, B = num_envs, NUM_DOF, NUM_BODIES
    rough = task == "pointfoot_rough"
    g = np.random.default_rng(seed * 1000003 + env_id_offset)
    f32 = np.float32
    s = {}

    root = np.zeros((N, 13), f32)
    if rough:
        root[:, 0] = g.uniform(0, 80, N)
        root[:, 1] = g.uniform(0, 160, N)
    else:   # env grid, spacing 3 m (point_foot.py:989-997) + small offset
        nc = np.floor(np.sqrt(N)); gid = np.arange(N)
        root[:, 0] = 3.0 * (gid // nc) + g.normal(0, 0.3, N)
        root[:, 1] = 3.0 * (gid % nc) + g.normal(0, 0.3, N)
    root[:, 2] = 0.62 + g.normal(0, 0.05, N)
    q = np.concatenate([g.normal(0, 0.15, (N, 3)), np.ones((N, 1))], 1)
    root[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    root[:, 7:13] = g.normal(0, 0.5, (N, 6))
    s["root_states"] = root

    dof = np.zeros((N, D, 2), f32)
    dof[..., 0] = g.uniform(DOF_LOWER, DOF_UPPER, (N, D))
    dof[..., 1] = g.normal(0, 3, (N, D))
    s["dof_state"] = dof.reshape(N * D, 2)

    cf = np.zeros((N, B, 3), f32)
    for f in FEET:
        cf[:, f, 2] = np.maximum(0, g.normal(40, 60, N))
        cf[:, f, :2] = g.normal(0, 5, (N, 2))
    hit = g.random((N, B)) < 0.002
    hit[:, list(FEET)] = False
    d = g.normal(0, 1, (N, B, 3)); d /= np.linalg.norm(d, axis=-1, keepdims=True)
    cf += (hit[..., None] * d * 5.0).astype(f32)
    s["contact_forces"] = cf.reshape(N * B, 3)

    rb = g.normal(0, 1, (N, B, 13)).astype(f32)
    for k, f in enumerate(FEET):
        rb[:, f, 0] = root[:, 0] + g.normal(0, 0.05, N)
        rb[:, f, 1] = root[:, 1] + (0.1 if k == 0 else -0.1) + g.normal(0, 0.05, N)
        rb[:, f, 2] = root[:, 2] - 0.55 + g.normal(0, 0.05, N)
    s["rigid_body_states"] = rb.reshape(N * B, 13)

    C = 4 if rough else 3
    cmd = np.zeros((N, C), f32)
    cmd[:, 0] = g.uniform(-1, 1, N); cmd[:, 1] = g.uniform(-0.2, 0.2, N)
    cmd[:, 2] = g.uniform(-1, 1, N) if rough else g.uniform(-1.5, 1.5, N)
    if rough:
        cmd[:, 3] = g.uniform(-3.14, 3.14, N)
    s["commands"] = cmd
    ep = g.integers(0, 1000, N)
    late = g.random(N) < 1e-3
    ep[late] = g.integers(1000, 1002, late.sum())
    s["episode_length_buf"] = ep.astype(np.int64)
    s["actions"] = g.normal(0, 1, (N, D)).astype(f32)
    s["base_mass"] = (13.0 + g.uniform(-1, 2, N)).astype(f32)

    # env-owned running state (randomised so that every branch of the feet state machine is exercised)
    s["last_actions"] = g.normal(0, 1, (N, D)).astype(f32)
    s["last_dof_vel"] = g.normal(0, 3, (N, D)).astype(f32)
    s["feet_air_time"] = (g.uniform(0, 0.8, (N, 2)) * (g.random((N, 2)) < 0.7)).astype(f32)
    s["last_feet_air_time"] = g.uniform(0, 0.8, (N, 2)).astype(f32)
    s["current_max_feet_height"] = g.uniform(0, 0.2, (N, 2)).astype(f32)
    s["last_max_feet_height"] = g.uniform(0, 0.2, (N, 2)).astype(f32)
    s["contact_filt"] = g.random((N, 2)) < 0.5
    s["first_contact"] = g.random((N, 2)) < 0.2
    s["last_contacts"] = g.random((N, 2)) < 0.5
    s["episode_sums"] = g.normal(0, 1, (num_reward_terms, N)).astype(f32)

    if rough:
        hf, origins = terrain if terrain is not None else make_heightfield(seed)
        s["height_samples"], s["terrain_origins"] = hf, origins
        s["terrain_levels"] = g.integers(0, 6, N).astype(np.int64)
        # point_foot.py:981-983 with the GLOBAL env count unknown to a shard: callers that shard pass
        # terrain_types themselves; single-shard default follows the reference formula.
        s["terrain_types"] = np.floor(np.arange(N) / (N / TERRAIN_COLS)).astype(np.int64)

    if adversarial:
        _inject_adversarial(s, N, rough)
    out = {k: (torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) else v)
           for k, v in s.items()}
    out["common_step_counter"] = 0
    return out


def _inject_adversarial(s, N, rough):
    """Exact-threshold and boundary cases: ||F||=1.0 exactly, Fz=1.0 / 0.1 exactly,
    episode counters at 1000/1001, positions outside the heightfield, resample-period multiples."""
    cf = s["contact_forces"].reshape(N, NUM_BODIES, 3)
    n = min(N, 16)
    for i in range(n):
        b = (1, 5, 0)[i % 3]
        cf[i, b] = 0
        cf[i, b, i % 3] = (1.0, np.nextafter(np.float32(1.0), np.float32(2.0)), 0.1, 0.10000001)[i % 4]
        cf[i, FEET[i % 2], 2] = (1.0, 0.1, np.nextafter(np.float32(1.0), np.float32(2.0)), 0.0)[(i // 2) % 4]
    s["episode_length_buf"][:n] = np.array([999, 1000, 1001, 199, 499, 0, 200, 500] * 2)[:n]
    if rough:
        r = s["root_states"]
        r[:n, 0] = np.array([-30.0, -25.0, -24.95, 104.9, 105.0, 300.0, 0.0, 79.95] * 2)[:n]
        r[:n, 1] = np.array([-30.0, 184.9, 185.0, -24.99, 500.0, 0.05, 0.1, 159.9] * 2)[:n]


# ------------------------------------------------------------------------------------------------ LeggedRobot / a1 (N2)
LR_NUM_DOF, LR_NUM_BODIES, LR_FEET = 12, 17, (4, 8, 12, 16)
LR_DOF_LOWER = np.array([-0.802851455917, -1.0471975512, -2.69653369433] * 4, np.float32)      # a1.urdf joint limits
LR_DOF_UPPER = np.array([0.802851455917, 4.18879020479, -0.916297857297] * 4, np.float32)


def make_lr_state(num_envs, seed=1, env_id_offset=0, adversarial=False, num_reward_terms=20, terrain=None):
    """Seeded synthetic state for the LeggedRobot base class with the a1 asset (12 DOF, 17 bodies, 4 feet) on the rough
    terrain (a1_config.py: trimesh, 187-point scan).  Same conventions as ``make_state``."""
    N, D, B = num_envs, LR_NUM_DOF, LR_NUM_BODIES
    g = np.random.default_rng(seed * 1000003 + env_id_offset + 77)
    f32 = np.float32
    s = {}
    root = np.zeros((N, 13), f32)
    root[:, 0] = g.uniform(0, 80, N)
    root[:, 1] = g.uniform(0, 160, N)
    root[:, 2] = 0.42 + g.normal(0, 0.05, N)
    q = np.concatenate([g.normal(0, 0.15, (N, 3)), np.ones((N, 1))], 1)
    root[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
    root[:, 7:13] = g.normal(0, 0.5, (N, 6))
    s["root_states"] = root
    dof = np.zeros((N, D, 2), f32)
    # inside the soft limits (a1_config.py: soft_dof_pos_limit 0.9) except ~1.5 % of the joints, which sit beyond them so that
    # dof_pos_limits (scale -10) fires for some envs and only_positive_rewards clips some totals and not others
    mid, span = (LR_DOF_UPPER + LR_DOF_LOWER) / 2, LR_DOF_UPPER - LR_DOF_LOWER
    u = g.uniform(-0.42, 0.42, (N, D))
    out_ = g.random((N, D)) < 0.015
    u = np.where(out_, np.sign(u) * g.uniform(0.455, 0.5, (N, D)), u)
    dof[..., 0] = mid + u * span
    dof[..., 1] = g.normal(0, 3, (N, D))
    # 40 % of the envs "track well" (joints near the default pose, slow joints, base velocity near the command): small torque /
    # acceleration penalties and large tracking rewards, so that the total is positive for them and the clip of
    # only_positive_rewards (LR:198-199) bites for the others
    good = g.random(N) < 0.4
    dflt = np.array([-0.1, 0.8, -1.5, 0.1, 0.8, -1.5, -0.1, 1.0, -1.5, 0.1, 1.0, -1.5], f32)     # a1_config.py default_joint_angles, DOF order
    dof[good, :, 0] = dflt + g.normal(0, 0.08, (int(good.sum()), D))
    dof[good, :, 1] = g.normal(0, 0.4, (int(good.sum()), D))
    s["_good"] = good
    s["dof_state"] = dof.reshape(N * D, 2)
    cf = np.zeros((N, B, 3), f32)
    for f in LR_FEET:
        cf[:, f, 2] = np.maximum(0, g.normal(20, 40, N))
        cf[:, f, :2] = g.normal(0, 5, (N, 2))
    hit = g.random((N, B)) < 0.004
    hit[:, list(LR_FEET)] = False
    d = g.normal(0, 1, (N, B, 3)); d /= np.linalg.norm(d, axis=-1, keepdims=True)
    cf += (hit[..., None] * d * 5.0).astype(f32)
    s["contact_forces"] = cf.reshape(N * B, 3)
    cmd = np.zeros((N, 4), f32)
    cmd[:, 0] = g.uniform(-1, 1, N); cmd[:, 1] = g.uniform(-1, 1, N)
    cmd[:, :2] *= (g.random((N, 1)) > 0.15)                   # some zero commands: feet_air_time's command gate
    cmd[:, 2] = g.uniform(-1, 1, N); cmd[:, 3] = g.uniform(-3.14, 3.14, N)
    s["commands"] = cmd
    ep = g.integers(0, 1000, N)
    late = g.random(N) < 1e-3
    ep[late] = g.integers(1000, 1002, late.sum())
    s["episode_length_buf"] = ep.astype(np.int64)
    s["actions"] = g.normal(0, 1, (N, D)).astype(f32)
    s["last_actions"] = g.normal(0, 1, (N, D)).astype(f32)
    s["last_dof_vel"] = g.normal(0, 3, (N, D)).astype(f32)
    good = s.pop("_good")
    ng = int(good.sum())
    s["last_dof_vel"][good] = dof[good, :, 1] + g.normal(0, 0.2, (ng, D)).astype(f32)
    s["last_actions"][good] = s["actions"][good] + g.normal(0, 0.2, (ng, D)).astype(f32)
    root[good, 7:9] = cmd[good, :2] + g.normal(0, 0.1, (ng, 2))          # (base frame ~ world frame: the quaternions are near identity)
    root[good, 9] = g.normal(0, 0.05, ng)
    root[good, 10:12] = g.normal(0, 0.1, (ng, 2))
    root[good, 12] = cmd[good, 2] + g.normal(0, 0.1, ng)
    s["feet_air_time"] = (g.uniform(0, 0.8, (N, 4)) * (g.random((N, 4)) < 0.7)).astype(f32)
    s["last_contacts"] = g.random((N, 4)) < 0.5
    s["episode_sums"] = g.normal(0, 1, (num_reward_terms, N)).astype(f32)
    hf, origins = terrain if terrain is not None else make_heightfield(seed)
    s["height_samples"], s["terrain_origins"] = hf, origins
    s["terrain_levels"] = g.integers(0, 6, N).astype(np.int64)
    s["terrain_types"] = np.floor(np.arange(N) / (N / TERRAIN_COLS)).astype(np.int64)
    if adversarial:
        n = min(N, 16)
        c = s["contact_forces"].reshape(N, B, 3)
        for i in range(n):
            c[i, 0] = 0
            c[i, 0, i % 3] = (1.0, np.nextafter(f32(1.0), f32(2.0)), 0.1, 0.10000001)[i % 4]      # termination body: ||F|| at 1.0
            c[i, LR_FEET[i % 4], 2] = (1.0, 0.1, np.nextafter(f32(1.0), f32(2.0)), 0.0)[(i // 2) % 4]
            c[i, 2 + 4 * (i % 4)] = 0
            c[i, 2 + 4 * (i % 4), 0] = (0.1, 0.10000001)[i % 2]                                  # penalised body at 0.1
        s["episode_length_buf"][:n] = np.array([999, 1000, 1001, 499, 500, 0, 1000, 500] * 2)[:n]   # resample period 500
        r = s["root_states"]
        r[:n, 0] = np.array([-30.0, -25.0, -24.95, 104.9, 105.0, 300.0, 0.0, 79.95] * 2)[:n]
        r[:n, 1] = np.array([-30.0, 184.9, 185.0, -24.99, 500.0, 0.05, 0.1, 159.9] * 2)[:n]
    out = {k: (torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) else v) for k, v in s.items()}
    out["common_step_counter"] = 0
    return out
