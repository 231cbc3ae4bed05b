"""cfg -> flat POD parameter block for the kernels.

Restates the host-side constant preparation the reference does at init time:
``_parse_cfg`` (point_foot.py:999-1009), ``_process_dof_props`` soft limits (:449-463), body index lookup (:910-968),
PD gains (:777-794), ``_prepare_reward_function`` (:796-820), ``_get_noise_scale_vec`` (:652-695, including its
column misalignment) and ``_init_height_points`` (:1032-1046).
"""
import numpy as np
import torch

from . import capi
from .config import class_to_dict


class RobotModel:
    """What PhysX reports for the asset after fixed-joint collapse (PF_P441A.urdf:148,191,234,390,433,476)."""
    dof_names = ["abad_L_Joint", "hip_L_Joint", "knee_L_Joint", "abad_R_Joint", "hip_R_Joint", "knee_R_Joint"]
    body_names = ["base_Link", "abad_L_Link", "hip_L_Link", "knee_L_Link", "foot_L_Link",
                  "abad_R_Link", "hip_R_Link", "knee_R_Link", "foot_R_Link"]
    dof_lower = [-0.38397, -1.012291, -0.872665, -1.39626, -1.518436, -1.396263]
    dof_upper = [1.39626, 1.518436, 1.396263, 0.38397, 1.012291, 0.872665]
    dof_velocity = [20.0, 20.0, 20.0, 20.0, 20.0, 50.0]
    dof_effort = [80.0] * 6
    base_mass = 13.0


def body_indices(robot, names):
    out = []
    for n in names:
        out.extend(i for i, s in enumerate(robot.body_names) if n in s)
    return out


def soft_dof_pos_limits(robot, soft):
    """point_foot.py:455-463, evaluated in fp32 tensor arithmetic like the reference."""
    lim = torch.zeros(len(robot.dof_lower), 2, dtype=torch.float)
    for i in range(len(robot.dof_lower)):
        lim[i, 0] = float(np.float32(robot.dof_lower[i]))
        lim[i, 1] = float(np.float32(robot.dof_upper[i]))
        m = (lim[i, 0] + lim[i, 1]) / 2
        r = lim[i, 1] - lim[i, 0]
        lim[i, 0] = m - 0.5 * r * soft
        lim[i, 1] = m + 0.5 * r * soft
    return lim


def reward_table(cfg, dt):
    """Non-zero scales * dt in ``dir()`` (alphabetical) order; 'termination' split off (point_foot.py:796-820)."""
    scales = class_to_dict(cfg.rewards.scales)
    names, vals, term = [], [], None
    for key in list(scales.keys()):
        s = scales[key]
        if s == 0:
            continue
        s = s * dt
        if key == "termination":
            term = s
            continue
        fn = key                       # '_reward_' + name must exist (getattr at point_foot.py:815)
        if fn not in capi.REWARD_IDS:
            raise AttributeError(f"'PointFoot' object has no attribute '_reward_{key}'")
        names.append(key)
        vals.append(s)
    return names, vals, term


def noise_scale_vec(cfg, num_dof=6):
    """point_foot.py:652-695.  Returns one fp32 vector of length max(num_obs, num_priv_obs): the obs part followed
    by the privileged-extra part.  (The LR column layout is applied to the PF obs layout exactly as the
    reference does.)"""
    n_obs = cfg.env.num_propriceptive_obs
    n_priv = cfg.env.num_privileged_obs
    ns, lvl, os_ = cfg.noise.noise_scales, cfg.noise.noise_level, cfg.normalization.obs_scales
    v = torch.zeros(n_obs)
    v[:3] = ns.lin_vel * lvl * os_.lin_vel
    v[3:6] = ns.ang_vel * lvl * os_.ang_vel
    v[6:9] = ns.gravity * lvl
    ce = 9 + cfg.commands.num_commands
    v[9:ce] = 0.
    pe = ce + num_dof
    v[ce:pe] = ns.dof_pos * lvl * os_.dof_pos
    ve = pe + num_dof
    v[pe:ve] = ns.dof_vel * lvl * os_.dof_vel
    ae = ve + cfg.env.num_actions
    v[ve:ae] = 0.
    npts = len(cfg.terrain.measured_points_x) * len(cfg.terrain.measured_points_y)
    extra = torch.zeros(n_priv - n_obs) if n_priv is not None else None
    if cfg.terrain.measure_heights_actor:
        v[ae:ae + npts] = ns.height_measurements * lvl * os_.height_measurements
    if cfg.terrain.measure_heights_critic and extra is not None:
        extra[:npts] = ns.height_measurements * lvl * os_.height_measurements
    return v, extra


def height_points_xy(cfg):
    """point_foot.py:1038-1045: meshgrid(x, y) ('ij'), flattened; fp32[P,2]."""
    x = torch.tensor(cfg.terrain.measured_points_x, dtype=torch.float)
    y = torch.tensor(cfg.terrain.measured_points_y, dtype=torch.float)
    gx, gy = torch.meshgrid(x, y, indexing="ij")
    return torch.stack((gx.flatten(), gy.flatten()), dim=1).contiguous()


def build_params(cfg, num_envs, robot=RobotModel, seed=1, env_id_offset=0, hf_shape=(0, 0),
                 index_div_mode=capi.DIV_TRUE, resample_commands=False):
    """Returns (PfParams, noise_vec fp32[max(num_obs,num_priv)], height_points fp32[P,2], reward_names)."""
    p = capi.PfParams()
    D = len(robot.dof_names)
    sim_dt = cfg.sim.dt
    dt = cfg.control.decimation * sim_dt
    p.abi_version = capi.PF_ABI_VERSION
    p.num_envs, p.env_id_offset = num_envs, env_id_offset
    p.num_dof, p.num_bodies = D, len(robot.body_names)
    p.num_commands = cfg.commands.num_commands
    p.num_obs = cfg.env.num_propriceptive_obs
    p.num_priv_obs = cfg.env.num_privileged_obs or 0
    p.seed = seed & 0xFFFFFFFF
    p.control_type = capi.CONTROL_TYPES.get(cfg.control.control_type, -1)
    if p.control_type < 0:
        raise NameError(f"Unknown controller type: {cfg.control.control_type}")      # point_foot.py:546
    p.decimation = cfg.control.decimation
    p.action_scale, p.sim_dt, p.dt = cfg.control.action_scale, sim_dt, dt
    p.clip_actions, p.clip_obs = cfg.normalization.clip_actions, cfg.normalization.clip_observations
    lim = soft_dof_pos_limits(robot, cfg.rewards.soft_dof_pos_limit)
    for i, name in enumerate(robot.dof_names):
        p.default_dof_pos[i] = cfg.init_state.default_joint_angles[name]
        kp = kd = 0.0
        for key in cfg.control.stiffness.keys():
            if key in name:
                kp, kd = cfg.control.stiffness[key], cfg.control.damping[key]
        p.p_gains[i], p.d_gains[i] = kp, kd
        p.torque_limits[i] = robot.dof_effort[i]
        p.dof_vel_limits[i] = robot.dof_velocity[i]
        p.dof_pos_limits[i][0], p.dof_pos_limits[i][1] = float(lim[i, 0]), float(lim[i, 1])
    feet = [i for i, s in enumerate(robot.body_names) if cfg.asset.foot_name in s]
    assert len(feet) == capi.NUM_FEET
    for i, b in enumerate(feet):
        p.feet_indices[i] = b
    term = body_indices(robot, cfg.asset.terminate_after_contacts_on)
    pen = body_indices(robot, cfg.asset.penalize_contacts_on)
    p.num_termination, p.num_penalised = len(term), len(pen)
    for i, b in enumerate(term):
        p.termination_indices[i] = b
    for i, b in enumerate(pen):
        p.penalised_indices[i] = b
    p.max_episode_length_s = cfg.env.episode_length_s
    p.max_episode_length = float(np.ceil(cfg.env.episode_length_s / dt))
    p.resample_period = int(cfg.commands.resampling_time / dt)
    p.heading_command = int(bool(cfg.commands.heading_command))
    p.resample_commands = int(bool(resample_commands))
    r = cfg.commands.ranges
    for i, rng in enumerate((r.lin_vel_x, r.lin_vel_y, r.ang_vel_yaw, r.heading)):
        p.cmd_range[i][0], p.cmd_range[i][1] = rng[0], rng[1]
    os_ = cfg.normalization.obs_scales
    p.obs_scale_ang_vel, p.obs_scale_dof_pos = os_.ang_vel, os_.dof_pos
    p.obs_scale_dof_vel, p.obs_scale_height = os_.dof_vel, os_.height_measurements
    p.commands_scale[0], p.commands_scale[1], p.commands_scale[2] = os_.lin_vel, os_.lin_vel, os_.ang_vel
    p.add_noise = int(bool(cfg.noise.add_noise))
    p.measure_heights_actor = int(bool(cfg.terrain.measure_heights_actor))
    p.measure_heights_critic = int(bool(cfg.terrain.measure_heights_critic))
    mesh = cfg.terrain.mesh_type
    if mesh not in (None, "plane", "heightfield", "trimesh"):
        raise ValueError("Terrain mesh type not recognised. Allowed types are [None, plane, heightfield, trimesh]")
    rough = mesh in ("heightfield", "trimesh")
    p.terrain_kind = capi.TERRAIN_HEIGHTFIELD if rough else capi.TERRAIN_PLANE
    p.index_div_mode = index_div_mode
    p.hf_rows, p.hf_cols = hf_shape
    p.border_size, p.horizontal_scale = cfg.terrain.border_size, cfg.terrain.horizontal_scale
    p.vertical_scale = cfg.terrain.vertical_scale
    measures = p.measure_heights_actor or p.measure_heights_critic
    pts = height_points_xy(cfg) if measures else torch.zeros(0, 2)
    if measures and not rough and mesh != "plane":
        raise NameError("Can't measure height with terrain mesh type 'none'")        # point_foot.py:1064
    p.num_height_points = pts.shape[0]
    if p.num_height_points > capi.MAX_HEIGHT_POINTS:
        raise ValueError("too many height points")
    p.custom_origins = int(rough)
    p.terrain_curriculum = int(bool(cfg.terrain.curriculum) and rough)            # point_foot.py:1004-1005
    p.max_terrain_level, p.num_terrain_types = cfg.terrain.num_rows, cfg.terrain.num_cols
    p.terrain_env_length = cfg.terrain.terrain_length
    ist = cfg.init_state
    for i, v in enumerate(ist.pos + ist.rot + ist.lin_vel + ist.ang_vel):
        p.base_init_state[i] = v
    p.send_timeouts = int(bool(cfg.env.send_timeouts))
    names, vals, term_scale = reward_table(cfg, dt)
    if len(names) > capi.MAX_REWARD_TERMS:
        raise ValueError("too many reward terms")
    p.num_reward_terms = len(names)
    for i, (n, s) in enumerate(zip(names, vals)):
        p.reward_ids[i], p.reward_scales[i] = capi.REWARD_IDS[n], s
    p.has_termination_term = int(term_scale is not None)
    p.termination_scale = term_scale or 0.0
    p.only_positive_rewards = int(bool(cfg.rewards.only_positive_rewards))
    all_scales = [a for a in dir(cfg.rewards.scales) if not a.startswith("__")]
    p.include_feet_height_rewards = int(any("feet_height" in a for a in all_scales))   # point_foot.py:232-237
    rw = cfg.rewards
    p.base_height_target, p.soft_dof_vel_limit = rw.base_height_target, rw.soft_dof_vel_limit
    p.soft_torque_limit, p.tracking_sigma = rw.soft_torque_limit, rw.tracking_sigma
    p.max_contact_force = rw.max_contact_force
    p.min_feet_air_time, p.max_feet_air_time = rw.min_feet_air_time, rw.max_feet_air_time
    p.min_feet_distance = rw.min_feet_distance
    p.max_push_vel_xy = cfg.domain_rand.max_push_vel_xy
    v, extra = noise_scale_vec(cfg, D)
    noise = torch.cat((v, extra)) if extra is not None else v
    if measures and rough is False and p.num_height_points:
        pass   # 'plane' + measuring: heights are zeros (point_foot.py:1061-1062); handled in the kernel
    if names and term_scale is not None:
        names = names + ["termination"]
    elif term_scale is not None:
        names = ["termination"]
    return p, noise.contiguous(), pts, names
