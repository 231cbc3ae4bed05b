"""``LeggedRobot`` -- drop-in for legged_gym/envs/base/legged_robot.py::LeggedRobot (LR) on the per-step path, with the a1
task (legged_gym/envs/a1/a1_config.py) as its instance.  SURVEY.md section 8f, row N2.

Same constructor signature, attributes and VecEnv protocol as the reference class (LR:48-70, 76-103; base_task.py:40-112);
every per-step computation is a launch into libpointfoot_b200.so through the context-free ``pf_lr_*`` entry points
(include/pointfoot_b200.h, "LeggedRobot").  Physics stays behind the ``gym`` object as in the reference (LR:91-97); the
default backend is the inert ``SyntheticGym`` over seeded tensors.  No torch arithmetic on the path, no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import capi, synth
from .config import class_to_dict
from .env import SyntheticGym, _Wrap
from .params import body_indices, height_points_xy, soft_dof_pos_limits


class A1Model:
    """What PhysX reports for resources/robots/a1/urdf/a1.urdf after fixed-joint collapse (the feet are kept by
    dont_collapse, a1.urdf:187,309,431,553): 12 revolute joints, 17 bodies."""
    dof_names = [f"{leg}_{j}_joint" for leg in ("FR", "FL", "RR", "RL") for j in ("hip", "thigh", "calf")]
    body_names = ["base"] + [f"{leg}_{b}" for leg in ("FR", "FL", "RR", "RL") for b in ("hip", "thigh", "calf", "foot")]
    dof_lower = [-0.802851455917, -1.0471975512, -2.69653369433] * 4
    dof_upper = [0.802851455917, 4.18879020479, -0.916297857297] * 4
    dof_velocity = [52.4, 28.6, 28.6] * 4
    dof_effort = [20.0, 55.0, 55.0] * 4


# the _reward_* methods LeggedRobot defines (LR:816-906); anything else fails like the reference's getattr (LR:551)
LR_REWARD_NAMES = ("lin_vel_z", "ang_vel_xy", "orientation", "base_height", "torques", "dof_vel", "dof_acc", "action_rate",
                   "collision", "termination", "dof_pos_limits", "dof_vel_limits", "torque_limits", "tracking_lin_vel",
                   "tracking_ang_vel", "feet_air_time", "stumble", "stand_still", "feet_contact_forces")


def lr_reward_table(cfg, dt):
    """LR:540-560: non-zero scales * dt in ``dir()`` order, 'termination' split off (it is added after the clip, LR:203-206)."""
    names, vals, term = [], [], None
    for key, s in class_to_dict(cfg.rewards.scales).items():
        if s == 0:
            continue
        if key == "termination":
            term = s * dt
            continue
        if key not in LR_REWARD_NAMES:
            raise AttributeError(f"'LeggedRobot' object has no attribute '_reward_{key}'")
        names.append(key)
        vals.append(s * dt)
    return names, vals, term


def lr_noise_scale_vec(cfg):
    """LR:455-478, including its hard-coded column numbers."""
    ns, lvl, os_ = cfg.noise.noise_scales, cfg.noise.noise_level, cfg.normalization.obs_scales
    v = torch.zeros(cfg.env.num_observations)
    v[:3] = ns.lin_vel * lvl * os_.lin_vel
    v[3:6] = ns.ang_vel * lvl * os_.ang_vel
    v[6:9] = ns.gravity * lvl
    v[12:24] = ns.dof_pos * lvl * os_.dof_pos
    v[24:36] = ns.dof_vel * lvl * os_.dof_vel
    if cfg.terrain.measure_heights:
        v[48:235] = ns.height_measurements * lvl * os_.height_measurements
    return v


def build_lr_params(cfg, num_envs, robot=A1Model, seed=1, env_id_offset=0, hf_shape=(0, 0), index_div_mode=capi.DIV_TRUE):
    """cfg -> pf_lr_params: _parse_cfg (LR:728-740), _process_dof_props (LR:279-305), _create_envs index look-ups
    (LR:680-699), PD gains (LR:523-539), _prepare_reward_function (LR:540-560).
    Returns (PfLrParams, noise fp32[num_obs], height points fp32[P,2], reward row names with 'termination' last)."""
    p = capi.PfLrParams()
    D = len(robot.dof_names)
    if D > capi.LR_MAX_DOF or len(robot.body_names) > capi.LR_MAX_BODIES:
        raise ValueError("asset larger than PF_LR_MAX_DOF / PF_LR_MAX_BODIES")
    sim_dt = cfg.sim.dt
    dt = cfg.control.decimation * sim_dt
    p.abi_version, p.num_envs, p.env_id_offset = capi.PF_ABI_VERSION, num_envs, env_id_offset
    p.num_dof, p.num_bodies, p.num_commands = D, len(robot.body_names), cfg.commands.num_commands
    p.num_obs, p.seed = cfg.env.num_observations, seed & 0xFFFFFFFF
    p.control_type = capi.CONTROL_TYPES.get(cfg.control.control_type, -1)
    if p.control_type < 0:
        raise NameError(f"Unknown controller type: {cfg.control.control_type}")      # LR:374
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
        p.torque_limits[i], p.dof_vel_limits[i] = robot.dof_effort[i], robot.dof_velocity[i]
        p.dof_pos_limits[i][0], p.dof_pos_limits[i][1] = float(lim[i, 0]), float(lim[i, 1])
    feet = [i for i, s in enumerate(robot.body_names) if cfg.asset.foot_name in s]
    if len(feet) > capi.LR_MAX_FEET:
        raise ValueError("more feet than PF_LR_MAX_FEET")
    p.num_feet = len(feet)
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
    r = cfg.commands.ranges
    for i, rng in enumerate((r.lin_vel_x, r.lin_vel_y, r.ang_vel_yaw, r.heading)):
        p.cmd_range[i][0], p.cmd_range[i][1] = rng[0], rng[1]
    os_ = cfg.normalization.obs_scales
    p.obs_scale_lin_vel, p.obs_scale_ang_vel, p.obs_scale_dof_pos = os_.lin_vel, os_.ang_vel, os_.dof_pos
    p.obs_scale_dof_vel, p.obs_scale_height = os_.dof_vel, os_.height_measurements
    p.commands_scale[0], p.commands_scale[1], p.commands_scale[2] = os_.lin_vel, os_.lin_vel, os_.ang_vel   # LR:507
    p.add_noise, p.measure_heights = int(bool(cfg.noise.add_noise)), int(bool(cfg.terrain.measure_heights))
    mesh = cfg.terrain.mesh_type
    if mesh not in (None, "plane", "heightfield", "trimesh"):
        raise ValueError("Terrain mesh type not recognised. Allowed types are [None, plane, heightfield, trimesh]")   # LR:241
    rough = mesh in ("heightfield", "trimesh")
    if p.measure_heights and mesh is None:
        raise NameError("Can't measure height with terrain mesh type 'none'")        # LR:791
    p.terrain_kind = capi.TERRAIN_HEIGHTFIELD if rough else capi.TERRAIN_PLANE
    p.index_div_mode = index_div_mode
    p.hf_rows, p.hf_cols = hf_shape
    p.border_size, p.horizontal_scale, p.vertical_scale = cfg.terrain.border_size, cfg.terrain.horizontal_scale, cfg.terrain.vertical_scale
    pts = height_points_xy(cfg) if p.measure_heights else torch.zeros(0, 2)
    p.num_height_points = pts.shape[0]
    p.custom_origins = int(rough)
    p.terrain_curriculum = int(bool(cfg.terrain.curriculum) and rough)              # LR:733-734
    p.max_terrain_level, p.num_terrain_types = cfg.terrain.num_rows, cfg.terrain.num_cols
    p.terrain_env_length = cfg.terrain.terrain_length
    ist = cfg.init_state
    for i, v in enumerate(ist.pos + ist.rot + ist.lin_vel + ist.ang_vel):
        p.base_init_state[i] = v
    p.send_timeouts = int(bool(cfg.env.send_timeouts))
    names, vals, term_scale = lr_reward_table(cfg, dt)
    if len(names) + (term_scale is not None) > capi.MAX_REWARD_TERMS:
        raise ValueError("too many reward terms")
    p.num_reward_terms = len(names)
    for i, (n, s) in enumerate(zip(names, vals)):
        p.reward_ids[i], p.reward_scales[i] = capi.REWARD_IDS[n], s
    p.has_termination_term, p.termination_scale = int(term_scale is not None), term_scale or 0.0
    p.only_positive_rewards = int(bool(cfg.rewards.only_positive_rewards))
    rw = cfg.rewards
    p.base_height_target, p.soft_dof_vel_limit, p.soft_torque_limit = rw.base_height_target, rw.soft_dof_vel_limit, rw.soft_torque_limit
    p.tracking_sigma, p.max_contact_force = rw.tracking_sigma, rw.max_contact_force
    p.max_push_vel_xy = cfg.domain_rand.max_push_vel_xy
    if term_scale is not None:
        names = names + ["termination"]
    return p, lr_noise_scale_vec(cfg).contiguous(), pts.contiguous(), names


class LeggedRobot:
    """VecEnv over the ``pf_lr_*`` kernels.  ``robot`` is the asset description PhysX would report (default: a1)."""

    def __init__(self, cfg, sim_params=None, physics_engine=None, sim_device="cuda:0", headless=True,
                 gym=None, gymtorch=None, state=None, seed=1, env_id_offset=0, robot=A1Model):
        self.cfg, self.sim_params, self.physics_engine, self.headless = cfg, sim_params, physics_engine, headless
        self.viewer, self.enable_viewer_sync, self.debug_viz, self.init_done = None, True, False, False
        self.device, self.sim_device = torch.device(sim_device), str(sim_device)
        self._open_backend()
        self.env_id_offset, self.seed = int(env_id_offset), int(seed)
        N = self.num_envs = cfg.env.num_envs
        self.num_obs, self.num_privileged_obs, self.num_actions = cfg.env.num_observations, cfg.env.num_privileged_obs, cfg.env.num_actions
        if cfg.terrain.mesh_type not in ("heightfield", "trimesh"):
            cfg.terrain.curriculum = False                                           # LR:733-734
        rough = cfg.terrain.mesh_type in ("heightfield", "trimesh")
        if gym is None:
            if state is None:
                state = synth.make_lr_state(N, seed=seed, env_id_offset=env_id_offset)
            state = {k: (v.to(self.device).contiguous() if torch.is_tensor(v) else v) for k, v in state.items()}
            gym, gymtorch = SyntheticGym(state), _Wrap
        self.gym, self._gymtorch, self.sim = gym, gymtorch or _Wrap, getattr(gym, "sim", None)
        self._state = state or {}
        self.envs, self.actor_handles = [None] * N, [None] * N
        hf = self._state.get("height_samples") if rough else None
        self.height_samples = hf
        P, noise, pts, names = build_lr_params(cfg, N, robot, seed=seed, env_id_offset=env_id_offset,
                                               hf_shape=tuple(hf.shape) if hf is not None else (0, 0))
        self._P, self.reward_names = P, names
        self.noise_scale_vec, self._height_points_xy = noise.to(self.device), pts.to(self.device)
        self.add_noise = bool(P.add_noise)
        self.dt = cfg.control.decimation * cfg.sim.dt
        self.obs_scales = cfg.normalization.obs_scales
        self.reward_scales = {n: float(P.reward_scales[i]) for i, n in enumerate(names[:P.num_reward_terms])}
        self.command_ranges = class_to_dict(cfg.commands.ranges)
        self.max_episode_length_s = cfg.env.episode_length_s
        self.max_episode_length = np.ceil(self.max_episode_length_s / self.dt)
        cfg.domain_rand.push_interval = np.ceil(cfg.domain_rand.push_interval_s / self.dt)   # LR:738
        self.num_dof = self.num_dofs = P.num_dof
        self.num_bodies, self.dof_names = P.num_bodies, list(robot.dof_names)
        self.custom_origins = bool(P.custom_origins)
        self._init_buffers(names)
        self.init_done = True

    def _open_backend(self):
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("pointfoot_b200.LeggedRobot needs a CUDA device (no CPU fallback); got " + self.sim_device)
        self._lib = capi.load_lr()

    # the three C-ABI calls of the path (include/pointfoot_b200.h, "LeggedRobot")
    def _call_pd_torques(self, actions):
        capi.check(self._lib.pf_lr_pd_torques(*self._args(), None if actions is None else C.c_void_p(actions.data_ptr()),
                                              self._stream()), "pf_lr_pd_torques")

    def _call_post_physics_step(self, push):
        capi.check(self._lib.pf_lr_post_physics_step(*self._args(), C.c_int64(self.common_step_counter), C.c_int32(push),
                                                     self._stream()), "pf_lr_post_physics_step")

    def _call_reset_all(self):
        capi.check(self._lib.pf_lr_reset_all(*self._args(), C.c_int64(self.common_step_counter), self._stream()), "pf_lr_reset_all")

    # ------------------------------------------------------------------ buffers (base_task.py:65-80, LR:480-538, 701-726)
    def _init_buffers(self, names):
        P, N, dev = self._P, self.num_envs, self.device
        D, F = P.num_dof, P.num_feet
        z = lambda *s, dt=torch.float: torch.zeros(*s, dtype=dt, device=dev)
        vec = lambda a, n: torch.tensor([a[i] for i in range(n)], device=dev)
        wrap = self._gymtorch.wrap_tensor
        self.root_states = wrap(self.gym.acquire_actor_root_state_tensor(self.sim))
        self.dof_state = wrap(self.gym.acquire_dof_state_tensor(self.sim))
        self.dof_pos, self.dof_vel = self.dof_state.view(N, D, 2)[..., 0], self.dof_state.view(N, D, 2)[..., 1]
        self.base_quat = self.root_states[:, 3:7]
        self.contact_forces = wrap(self.gym.acquire_net_contact_force_tensor(self.sim)).view(N, -1, 3)
        self.feet_indices = vec(P.feet_indices, F).long()
        self.termination_contact_indices = vec(P.termination_indices, P.num_termination).long()
        self.penalised_contact_indices = vec(P.penalised_indices, P.num_penalised).long()
        self.common_step_counter = 0
        self.extras = {}
        self.obs_buf, self.privileged_obs_buf = z(N, self.num_obs), None            # (LeggedRobot never fills a privileged buffer)
        self.rew_buf = z(N)
        self.reset_buf = torch.ones(N, dtype=torch.bool, device=dev)                # long -> bool at LR:141
        self.episode_length_buf = z(N, dt=torch.long)
        self.time_out_buf = z(N, dt=torch.bool)
        self.torques, self.actions, self.last_actions = z(N, D), z(N, D), z(N, D)
        self.last_dof_vel, self.last_root_vel = z(N, D), z(N, 6)
        self.commands = z(N, P.num_commands)
        self.commands_scale = vec(P.commands_scale, 3)
        self.feet_air_time, self.last_contacts = z(N, F), z(N, F, dt=torch.bool)
        for k in ("commands", "episode_length_buf", "last_actions", "last_dof_vel", "feet_air_time", "last_contacts"):
            if k in self._state:                                                     # a synthetic mid-rollout state
                getattr(self, k).copy_(self._state[k])
        self.base_lin_vel, self.base_ang_vel, self.projected_gravity = z(N, 3), z(N, 3), z(N, 3)
        self.default_dof_pos = vec(P.default_dof_pos, D).unsqueeze(0)
        self.p_gains, self.d_gains, self.torque_limits = vec(P.p_gains, D), vec(P.d_gains, D), vec(P.torque_limits, D)
        self.measured_heights = z(N, P.num_height_points) if P.measure_heights else 0    # LR:521
        R = P.num_reward_terms + P.has_termination_term
        self._episode_sums = z(R, N)
        if "episode_sums" in self._state:
            self._episode_sums.copy_(self._state["episode_sums"][:R])
        self.episode_sums = {n: self._episode_sums[i] for i, n in enumerate(names)}
        self._episode_stats = z(R + 2)
        self._reset_ids, self._reset_count = z(N, dt=torch.int32), z(1, dt=torch.int32)
        self._time_outs = z(N, dt=torch.bool)
        if self.custom_origins:                                                       # LR:701-716
            self.terrain_origins = self._state["terrain_origins"]
            self.terrain_levels = self._state["terrain_levels"].clone()
            self.terrain_types = self._state["terrain_types"]
            self.max_terrain_level = P.max_terrain_level
            self.env_origins = self.terrain_origins[self.terrain_levels, self.terrain_types].clone().contiguous()
        else:                                                                         # LR:717-726
            self.terrain_origins = self.terrain_levels = self.terrain_types = None
            o = z(N, 3)
            num_cols = np.floor(np.sqrt(N)); num_rows = np.ceil(N / num_cols)
            xx, yy = torch.meshgrid(torch.arange(num_rows), torch.arange(num_cols), indexing="ij")
            sp = self.cfg.env.env_spacing
            o[:, 0] = (sp * xx.flatten()[:N]).to(dev); o[:, 1] = (sp * yy.flatten()[:N]).to(dev)
            self.env_origins = o
        fields = dict(
            root_states=self.root_states, dof_state=self.dof_state, contact_forces=self.contact_forces, actions=self.actions,
            torques=self.torques, commands=self.commands, episode_length_buf=self.episode_length_buf,
            base_lin_vel=self.base_lin_vel, base_ang_vel=self.base_ang_vel, projected_gravity=self.projected_gravity,
            last_actions=self.last_actions, last_dof_vel=self.last_dof_vel, last_root_vel=self.last_root_vel,
            feet_air_time=self.feet_air_time, last_contacts=self.last_contacts, episode_sums=self._episode_sums,
            rew_buf=self.rew_buf, reset_buf=self.reset_buf, time_out_buf=self.time_out_buf, obs_buf=self.obs_buf,
            measured_heights=self.measured_heights if torch.is_tensor(self.measured_heights) else None,
            height_samples=self.height_samples, terrain_origins=self.terrain_origins, terrain_levels=self.terrain_levels,
            terrain_types=self.terrain_types, env_origins=self.env_origins, reset_ids=self._reset_ids,
            reset_count=self._reset_count, episode_stats=self._episode_stats, extras_time_outs=self._time_outs,
            noise_scale_vec=self.noise_scale_vec, height_points_xy=self._height_points_xy if P.num_height_points else None)
        b = capi.PfLrBuffers()
        for name in capi.LR_BUFFER_FIELDS:
            t = fields[name]
            if t is not None and not t.is_contiguous():
                raise RuntimeError(f"buffer {name} must be contiguous")
            setattr(b, name, None if t is None else t.data_ptr())
        self._bufs, self._buf_refs = b, fields
        # extras: 0-d views the finalize kernel refreshes in place on steps with >= 1 reset (LR:148-149, 175-187)
        ep = {"rew_" + n: self._episode_stats[i] for i, n in enumerate(names)}
        if P.terrain_curriculum:
            ep["terrain_level"] = self._episode_stats[R]
        if self.cfg.commands.curriculum:
            ep["max_command_x"] = self.command_ranges["lin_vel_x"][1]
        self.extras["episode"] = ep
        if self.cfg.env.send_timeouts:
            self.extras["time_outs"] = self._time_outs

    # ------------------------------------------------------------------ VecEnv protocol (base_task.py:87-112)
    def get_observations(self):
        return self.obs_buf

    def get_privileged_observations(self):
        return self.privileged_obs_buf

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _args(self):
        return C.byref(self._P), C.byref(self._bufs)

    def reset(self):
        """base_task.py:105-109: reset_idx over every env, then one zero-action step."""
        self._call_reset_all()
        self._notify_gym_of_resets()
        obs, priv, _, _, _ = self.step(torch.zeros(self.num_envs, self.num_actions, device=self.device))
        return obs, priv

    def step(self, actions):
        """LR:76-103."""
        gym, sim = self.gym, self.sim
        a = actions.to(self.device, torch.float).contiguous()
        self.render()
        for i in range(self.cfg.control.decimation):
            self._call_pd_torques(a if i == 0 else None)
            gym.set_dof_actuation_force_tensor(sim, self._gymtorch.unwrap_tensor(self.torques))
            gym.simulate(sim)
            gym.fetch_results(sim, True)
            gym.refresh_dof_state_tensor(sim)
        self.post_physics_step()
        return self.obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf, self.extras

    def post_physics_step(self):
        """LR:105-136: at most four launches (scan, step, height columns, finalize)."""
        gym, sim = self.gym, self.sim
        gym.refresh_actor_root_state_tensor(sim)
        gym.refresh_net_contact_force_tensor(sim)
        self.common_step_counter += 1
        push = int(bool(self.cfg.domain_rand.push_robots) and self.common_step_counter % self.cfg.domain_rand.push_interval == 0)
        self._call_post_physics_step(push)
        if push and hasattr(gym, "set_actor_root_state_tensor"):                      # LR:419
            gym.set_actor_root_state_tensor(sim, self._gymtorch.unwrap_tensor(self.root_states))
        self._notify_gym_of_resets()
        if self.cfg.commands.curriculum and self.common_step_counter % self.max_episode_length == 0 \
                and int(self._reset_count.item()) > 0:                                # LR:160-161
            self.update_command_curriculum()

    def _notify_gym_of_resets(self):
        """LR:386-390, 408-412: the indexed setters want an int32 id list and a HOST count; the rows themselves were
        already written in place by the kernel.  The synthetic backend needs neither."""
        if not getattr(self.gym, "needs_host_reset_count", True):
            return
        n = int(self._reset_count.item())
        if n == 0:
            return
        ids = self._gymtorch.unwrap_tensor(self._reset_ids)
        self.gym.set_dof_state_tensor_indexed(self.sim, self._gymtorch.unwrap_tensor(self.dof_state), ids, n)
        self.gym.set_actor_root_state_tensor_indexed(self.sim, self._gymtorch.unwrap_tensor(self.root_states), ids, n)

    def update_command_curriculum(self, env_ids=None):
        """LR:443-453: widen lin_vel_x when the reset envs tracked well (the one host-side read of the step path, every
        max_episode_length steps)."""
        n = int(self._reset_count.item())
        if n == 0 or "tracking_lin_vel" not in self.reward_scales:
            return
        # (episode sums of the reset envs were already zeroed by the finalize kernel; its mean is what extras holds.  The
        # reference widens the range BEFORE the same reset_idx resamples, LR:160-166; here the new range applies from the
        # next step on -- once per max_episode_length steps, on the reset envs' first command only)
        mean_per_s = float(self.extras["episode"]["rew_tracking_lin_vel"])
        if mean_per_s * self.max_episode_length_s / self.max_episode_length > 0.8 * self.reward_scales["tracking_lin_vel"]:
            r, mc = self.command_ranges["lin_vel_x"], self.cfg.commands.max_curriculum
            r[0], r[1] = float(np.clip(r[0] - 0.5, -mc, 0.)), float(np.clip(r[1] + 0.5, 0., mc))
            self._P.cmd_range[0][0], self._P.cmd_range[0][1] = r[0], r[1]

    def render(self, sync_frame_time=True):
        return None                              # viewer: unavailable on this box

    def close(self):
        self._bufs = self._buf_refs = None
