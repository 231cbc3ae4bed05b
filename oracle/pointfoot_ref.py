"""Plain-torch CPU restatement of ``PointFoot.step`` and everything it calls.  ORACLE / TEST CODE.

Follows legged_gym/envs/pointfoot/point_foot.py (PF) op for op so that fp32 rounding is the reference's:
step PF:170-193, post_physics_step PF:195-227, check_termination PF:239-247, reset_idx PF:249-299,
compute_reward PF:301-318, observations PF:320-382, callback PF:479-493, torques PF:523-547,
resets PF:549-588, push PF:590-611, curriculum PF:613-636, heights PF:1048-1112, feet PF:1114-1129,
rewards PF:1132-1253.  Pinned against the unmodified reference by tests/golden (generated with
oracle/ref_harness.py where /root/reference exists).  Random draws use the contract in oracle/rng.py.
It takes the same flat parameter block (pointfoot_b200.capi.PfParams) the kernels take, so that the cfg
flattening in pointfoot_b200/params.py is covered by the golden tests as well.
"""
import numpy as np
import torch

from . import rng
from .isaacgym_torch_utils import quat_rotate_inverse, quat_rotate, quat_apply, normalize

REWARD_NAMES = (
    "lin_vel_z", "ang_vel_xy", "orientation", "base_height", "torques", "dof_vel", "dof_acc", "action_rate",
    "collision", "termination", "dof_pos_limits", "dof_vel_limits", "torque_limits", "tracking_lin_vel",
    "tracking_ang_vel", "feet_air_time", "no_fly", "unbalance_feet_air_time", "unbalance_feet_height",
    "stumble", "stand_still", "feet_contact_forces", "feet_distance", "survival")


def _f(x):
    """python float as the reference's cfg values are (python doubles; torch casts them per op)."""
    return float(x)


def grid_env_origins(num_envs, spacing):
    """PF:988-997: robots on a grid when there is no terrain."""
    o = torch.zeros(num_envs, 3)
    num_cols = np.floor(np.sqrt(num_envs))
    num_rows = np.ceil(num_envs / num_cols)
    xx, yy = torch.meshgrid(torch.arange(num_rows), torch.arange(num_cols), indexing="ij")
    o[:, 0] = spacing * xx.flatten()[:num_envs]
    o[:, 1] = spacing * yy.flatten()[:num_envs]
    return o


class RefEnv:
    """State + step of the restated env.  ``state`` as made by pointfoot_b200.synth.make_state; the four
    PhysX-owned tensors are used in place."""

    def __init__(self, params, noise_vec, height_points_xy, state):
        P = self.P = params
        N, D, B = P.num_envs, P.num_dof, P.num_bodies
        self.N, self.D = N, D
        f = lambda a, n: torch.tensor([a[i] for i in range(n)], dtype=torch.float)
        self.p_gains, self.d_gains = f(P.p_gains, D), f(P.d_gains, D)
        self.torque_limits, self.dof_vel_limits = f(P.torque_limits, D), f(P.dof_vel_limits, D)
        self.default_dof_pos = f(P.default_dof_pos, D).unsqueeze(0)
        self.dof_pos_limits = torch.tensor([[P.dof_pos_limits[i][0], P.dof_pos_limits[i][1]] for i in range(D)])
        self.feet_indices = torch.tensor([P.feet_indices[i] for i in range(2)])
        self.termination_idx = torch.tensor([P.termination_indices[i] for i in range(P.num_termination)])
        self.penalised_idx = torch.tensor([P.penalised_indices[i] for i in range(P.num_penalised)])
        self.commands_scale = f(P.commands_scale, 3)
        self.base_init_state = f(P.base_init_state, 13)
        self.noise_vec = noise_vec.clone()
        self.reward_names = [REWARD_NAMES[P.reward_ids[i]] for i in range(P.num_reward_terms)]
        self.reward_scales = [np.float32(P.reward_scales[i]).item() for i in range(P.num_reward_terms)]
        self.dt = np.float32(P.dt).item()          # NB: params hold fp32; see test_params for dt == 0.02 handling
        # PhysX-owned
        self.root_states = state["root_states"]
        self.dof_state = state["dof_state"]
        self.dof_pos = self.dof_state.view(N, D, 2)[..., 0]
        self.dof_vel = self.dof_state.view(N, D, 2)[..., 1]
        self.contact_forces = state["contact_forces"].view(N, B, 3)
        self.rigid_body_states = state["rigid_body_states"].view(N, B, 13)
        # env-owned
        z = lambda *s, dt=torch.float: torch.zeros(*s, dtype=dt)
        self.actions, self.torques = z(N, D), z(N, D)
        self.commands = state["commands"].clone()
        self.episode_length_buf = state["episode_length_buf"].clone()
        self.last_actions = state.get("last_actions", z(N, D)).clone()
        self.last_dof_vel = state.get("last_dof_vel", z(N, D)).clone()
        self.last_root_vel = z(N, 6)
        g = lambda k, d: state[k].clone() if k in state else d
        self.feet_air_time = g("feet_air_time", z(N, 2))
        self.last_feet_air_time = g("last_feet_air_time", z(N, 2))
        self.current_max_feet_height = g("current_max_feet_height", z(N, 2))
        self.last_max_feet_height = g("last_max_feet_height", z(N, 2))
        self.feet_height = z(N, 2)
        self.contact_filt = g("contact_filt", z(N, 2, dt=torch.bool))
        self.first_contact = g("first_contact", z(N, 2, dt=torch.bool))
        self.last_contacts = g("last_contacts", z(N, 2, dt=torch.bool))
        R = P.num_reward_terms + P.has_termination_term
        self.episode_sums = state["episode_sums"][:R].clone() if "episode_sums" in state else z(R, N)
        self.rew_buf = z(N)
        self.reset_buf = torch.ones(N, dtype=torch.bool)
        self.time_out_buf = z(N, dt=torch.bool)
        self.terminate = z(N, dt=torch.bool)
        self.base_lin_vel, self.base_ang_vel, self.projected_gravity = z(N, 3), z(N, 3), z(N, 3)
        self.gravity_vec = torch.tensor([0., 0., -1.]).repeat((N, 1))
        self.forward_vec = torch.tensor([1., 0., 0.]).repeat((N, 1))
        self.obs_buf = z(N, P.num_obs)
        self.priv_obs_buf = z(N, P.num_priv_obs) if P.num_priv_obs else None
        self.measured_heights = 0                                   # PF:775
        self.num_height_points = P.num_height_points
        if P.num_height_points:
            pts = z(N, P.num_height_points, 3)
            pts[:, :, 0] = height_points_xy[:, 0]
            pts[:, :, 1] = height_points_xy[:, 1]
            self.height_points = pts
        self.rough = P.terrain_kind == 1
        if self.rough:
            self.height_samples = state["height_samples"]
            self.terrain_origins = state["terrain_origins"]
            self.terrain_levels = state["terrain_levels"].clone()
            self.terrain_types = state["terrain_types"].clone()
            self.env_origins = self.terrain_origins[self.terrain_levels, self.terrain_types].clone()
        else:                                                       # grid, PF:988-997
            self.env_origins = state["env_origins"].clone() if "env_origins" in state else grid_env_origins(N, 3.0)
        self.base_mass = state["base_mass"].clone()
        self.common_step_counter = int(state.get("common_step_counter", 0))
        self.episode_stats = z(R + 2)
        self.reset_ids = torch.zeros(0, dtype=torch.int32)
        self.push_forces = z(N, B, 3)
        self.extras_time_outs = z(N, dt=torch.bool)
        # PF:492-493, 1009: 0 = never (the caller sets it from cfg.domain_rand; it is host-side state, not a kernel parameter)
        self.push_interval = 0
        self.sim_dt_py = None           # cfg.sim.dt as the python float the reference divides by (PF:595)
        self.max_push_vel_xy_py = None

    # -------------------------------------------------------------------------------- PF:523-547
    def compute_torques(self):
        P = self.P
        a = self.actions * _f(np.float32(P.action_scale))
        if P.control_type == 0:
            t = self.p_gains * (a + self.default_dof_pos - self.dof_pos) - self.d_gains * self.dof_vel
        elif P.control_type == 1:
            t = self.p_gains * (a - self.dof_vel) - self.d_gains * (self.dof_vel - self.last_dof_vel) / _f(np.float32(P.sim_dt))
        else:
            t = a
        return torch.clip(t, -self.torque_limits, self.torque_limits)

    # -------------------------------------------------------------------------------- PF:170-193
    def step(self, actions, physics=None):
        P = self.P
        self.actions = torch.clip(actions, -_f(P.clip_actions), _f(P.clip_actions))
        for _ in range(P.decimation):
            self.torques = self.compute_torques().view(self.torques.shape)
            if physics is not None:
                physics(self)
        self.post_physics_step()
        c = _f(P.clip_obs)
        self.obs_buf = torch.clip(self.obs_buf, -c, c)
        if self.priv_obs_buf is not None:
            self.priv_obs_buf = torch.clip(self.priv_obs_buf, -c, c)
        return self.obs_buf, self.priv_obs_buf, self.rew_buf, self.reset_buf

    # -------------------------------------------------------------------------------- PF:195-227
    def post_physics_step(self):
        P = self.P
        self.episode_length_buf += 1
        self.common_step_counter += 1
        q = self.root_states[:, 3:7]
        self.base_lin_vel[:] = quat_rotate_inverse(q, self.root_states[:, 7:10])
        self.base_ang_vel[:] = quat_rotate_inverse(q, self.root_states[:, 10:13])
        self.projected_gravity[:] = quat_rotate_inverse(q, self.gravity_vec)
        if P.measure_heights_actor or P.measure_heights_critic:
            self.measured_heights = self.get_heights()
        self.compute_feet_states()
        self.callback()
        self.check_termination()
        self.compute_reward()
        env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        self.reset_ids = env_ids.to(torch.int32)
        self.reset_idx(env_ids)
        self.compute_observations()
        self.last_actions[:] = self.actions[:]
        self.last_dof_vel[:] = self.dof_vel[:]
        self.last_root_vel[:] = self.root_states[:, 7:13]

    # -------------------------------------------------------------------------------- PF:1048-1112, math.py:38-42
    def terrain_heights(self, points):
        P = self.P
        points = points + _f(P.border_size)
        hs = _f(np.float32(P.horizontal_scale))
        if P.index_div_mode == 0:
            points = (points / hs).long()
        else:   # CUDA eager: multiply by the fp32 reciprocal
            points = (points * _f(np.float32(1.0) / np.float32(P.horizontal_scale))).long()
        px = torch.clip(points[:, :, 0].view(-1), 0, self.height_samples.shape[0] - 2)
        py = torch.clip(points[:, :, 1].view(-1), 0, self.height_samples.shape[1] - 2)
        h = torch.min(self.height_samples[px, py], self.height_samples[px + 1, py])
        return torch.min(h, self.height_samples[px, py + 1])

    def get_heights(self):
        if not self.rough:
            return torch.zeros(self.N, self.num_height_points)
        quat = self.root_states[:, 3:7].repeat(1, self.num_height_points)
        qy = quat.clone().view(-1, 4)
        qy[:, :2] = 0.
        qy = normalize(qy)
        pts = quat_apply(qy, self.height_points) + (self.root_states[:, :3]).unsqueeze(1)
        return self.terrain_heights(pts).view(self.N, -1) * _f(np.float32(self.P.vertical_scale))

    def heights_below_feet(self, feet_state):
        if not self.rough:
            return torch.zeros(self.N, 2)
        return self.terrain_heights(feet_state[:, :, :2]).view(self.N, -1) * _f(np.float32(self.P.vertical_scale))

    # -------------------------------------------------------------------------------- PF:1114-1129
    def compute_feet_states(self):
        self.feet_state = self.rigid_body_states[:, self.feet_indices, :]
        self.last_feet_air_time = self.feet_air_time * self.first_contact + self.last_feet_air_time * ~self.first_contact
        self.feet_air_time *= ~self.contact_filt
        if self.P.include_feet_height_rewards:
            self.last_max_feet_height = (self.current_max_feet_height * self.first_contact
                                         + self.last_max_feet_height * ~self.first_contact)
            self.current_max_feet_height *= ~self.contact_filt
            self.feet_height = self.feet_state[:, :, 2] - self.heights_below_feet(self.feet_state)
            self.current_max_feet_height = torch.max(self.current_max_feet_height, self.feet_height)
        contact = self.contact_forces[:, self.feet_indices, 2] > 1.
        self.contact_filt = torch.logical_or(contact, self.last_contacts)
        self.last_contacts = contact
        self.first_contact = (self.feet_air_time > 0.) * self.contact_filt
        self.feet_air_time += self.dt

    # -------------------------------------------------------------------------------- PF:479-493, LR:337-351
    def callback(self):
        P = self.P
        ids = (self.episode_length_buf % P.resample_period == 0).nonzero(as_tuple=False).flatten()
        if P.resample_commands and len(ids):
            ctx = rng.DrawContext(P.seed, self.common_step_counter, rng.RESAMPLE_CMD, ids.numpy(), P.env_id_offset)
            u = torch.from_numpy(ctx.uniform((len(ids), 4)))
            r = [[np.float32(P.cmd_range[i][0]).item(), np.float32(P.cmd_range[i][1]).item()] for i in range(4)]
            self.commands[ids, 0] = (r[0][1] - r[0][0]) * u[:, 0] + r[0][0]
            self.commands[ids, 1] = (r[1][1] - r[1][0]) * u[:, 1] + r[1][0]
            if P.heading_command:
                self.commands[ids, 3] = (r[3][1] - r[3][0]) * u[:, 2] + r[3][0]     # third draw of LR:346-349 (heading OR yaw): lane 2
            else:
                self.commands[ids, 2] = (r[2][1] - r[2][0]) * u[:, 2] + r[2][0]
            self.commands[ids, :2] *= (torch.norm(self.commands[ids, :2], dim=1) > 0.2).unsqueeze(1)
        if P.heading_command:
            fwd = quat_apply(self.root_states[:, 3:7], self.forward_vec)
            heading = torch.atan2(fwd[:, 1], fwd[:, 0])
            ang = self.commands[:, 3] - heading
            ang %= 2 * np.pi
            ang -= 2 * np.pi * (ang > np.pi)
            self.commands[:, 2] = torch.clip(0.5 * ang, -1., 1.)
        if self.push_interval and self.common_step_counter % self.push_interval == 0:      # PF:492-493
            self.push_robots(self.base_mass.mean().item() * self.max_push_vel_xy_py / self.sim_dt_py)

    # -------------------------------------------------------------------------------- PF:590-611
    def push_robots(self, max_push_force):
        """``max_push_force`` = base_mass.mean().item() * max_push_vel_xy / sim_params.dt, a python double (PF:592-596)."""
        P = self.P
        m = max_push_force
        ctx = rng.DrawContext(P.seed, self.common_step_counter, rng.PUSH, np.arange(self.N), P.env_id_offset)
        f = (m - (-m)) * torch.from_numpy(ctx.uniform((self.N, 3))) + (-m)
        self.push_forces[:] = 0
        self.push_forces[:, 0, 0:3] = quat_rotate(self.root_states[:, 3:7], f)
        self.push_forces[:, 0, 2] *= 0.5
        return self.push_forces

    # -------------------------------------------------------------------------------- PF:239-247
    def check_termination(self):
        n = torch.norm(self.contact_forces[:, self.termination_idx, :], dim=-1)
        self.reset_buf = torch.any(n > 1., dim=1)
        self.terminate = torch.any(n > 1., dim=1)
        self.time_out_buf = self.episode_length_buf > _f(np.float32(self.P.max_episode_length))
        self.reset_buf |= self.time_out_buf

    # -------------------------------------------------------------------------------- PF:301-318, 1132-1253
    def reward_term(self, name):
        P, cf = self.P, self.contact_forces
        g = lambda x: _f(np.float32(x))
        if name == "lin_vel_z":
            return torch.square(self.base_lin_vel[:, 2])
        if name == "ang_vel_xy":
            return torch.sum(torch.square(self.base_ang_vel[:, :2]), dim=1)
        if name == "orientation":
            return torch.sum(torch.square(self.projected_gravity[:, :2]), dim=1)
        if name == "base_height":
            bh = torch.mean(self.root_states[:, 2].unsqueeze(1) - self.measured_heights, dim=1)
            return torch.square(bh - g(P.base_height_target))
        if name == "torques":
            return torch.sum(torch.square(self.torques), dim=1)
        if name == "dof_vel":
            return torch.sum(torch.square(self.dof_vel), dim=1)
        if name == "dof_acc":
            return torch.sum(torch.square((self.last_dof_vel - self.dof_vel) / self.dt), dim=1)
        if name == "action_rate":
            return torch.sum(torch.square(self.last_actions - self.actions), dim=1)
        if name == "collision":
            return torch.sum(1. * (torch.norm(cf[:, self.penalised_idx, :], dim=-1) > 0.1), dim=1)
        if name == "termination":
            return self.reset_buf * ~self.time_out_buf
        if name == "dof_pos_limits":
            o = -(self.dof_pos - self.dof_pos_limits[:, 0]).clip(max=0.)
            o += (self.dof_pos - self.dof_pos_limits[:, 1]).clip(min=0.)
            return torch.sum(o, dim=1)
        if name == "dof_vel_limits":
            return torch.sum((torch.abs(self.dof_vel) - self.dof_vel_limits * g(P.soft_dof_vel_limit)).clip(min=0., max=1.), dim=1)
        if name == "torque_limits":
            return torch.sum((torch.abs(self.torques) - self.torque_limits * g(P.soft_torque_limit)).clip(min=0.), dim=1)
        if name == "tracking_lin_vel":
            err = torch.sum(torch.square(self.commands[:, :2] - self.base_lin_vel[:, :2]), dim=1)
            return torch.exp(-err / g(P.tracking_sigma))
        if name == "tracking_ang_vel":
            err = torch.square(self.commands[:, 2] - self.base_ang_vel[:, 2])
            return torch.exp(-err / g(P.tracking_sigma))
        if name == "feet_air_time":
            zero = torch.zeros_like(self.feet_air_time)
            lo = torch.sum(torch.min(self.feet_air_time - g(P.min_feet_air_time), zero) * self.first_contact, dim=1)
            hi = torch.sum(torch.min(g(P.max_feet_air_time) - self.feet_air_time, zero) * self.first_contact, dim=1)
            return lo + hi
        if name == "no_fly":
            contacts = cf[:, self.feet_indices, 2] > 0.1
            return 1. * (torch.sum(1. * contacts, dim=1) == 1)
        if name == "unbalance_feet_air_time":
            return torch.var(self.last_feet_air_time, dim=-1)
        if name == "unbalance_feet_height":
            return torch.var(self.last_max_feet_height, dim=-1)
        if name == "stumble":
            return torch.any(torch.norm(cf[:, self.feet_indices, :2], dim=2) > 5 * torch.abs(cf[:, self.feet_indices, 2]), dim=1)
        if name == "stand_still":
            lin = torch.abs(self.base_lin_vel[:, :2]) * (self.commands[:, :2] < 0.1)
            ang = (torch.abs(self.base_ang_vel[:, -1]) * (self.commands[:, 2] < 0.1)).unsqueeze(dim=-1)
            return torch.sum(torch.cat((lin, ang), dim=-1), dim=-1)
        if name == "feet_contact_forces":
            return torch.sum((torch.norm(cf[:, self.feet_indices, :], dim=-1) - g(P.max_contact_force)).clip(min=0.), dim=1)
        if name == "feet_distance":
            d = torch.norm(self.feet_state[:, 0, :2] - self.feet_state[:, 1, :2], dim=-1)
            return 0 + torch.clip(g(P.min_feet_distance) - d, 0, 1)
        if name == "survival":
            return (~self.reset_buf).float() * self.dt
        raise AttributeError(name)

    def compute_reward(self):
        P = self.P
        self.rew_buf[:] = 0.
        for k, name in enumerate(self.reward_names):
            rew = self.reward_term(name) * self.reward_scales[k]
            self.rew_buf += rew
            self.episode_sums[k] += rew
        if P.only_positive_rewards:
            self.rew_buf[:] = torch.clip(self.rew_buf[:], min=0.)
        if P.has_termination_term:
            rew = self.reward_term("termination") * _f(np.float32(P.termination_scale))
            self.rew_buf += rew
            self.episode_sums[P.num_reward_terms] += rew

    # -------------------------------------------------------------------------------- PF:249-299, 549-588, 613-636
    def reset_idx(self, env_ids, update_curriculum=True):
        P = self.P
        if len(env_ids) == 0:
            return
        ids_np = env_ids.numpy()
        step = self.common_step_counter
        if P.terrain_curriculum and update_curriculum:
            dist = torch.norm(self.root_states[env_ids, :2] - self.env_origins[env_ids, :2], dim=1)
            up = dist > _f(np.float32(P.terrain_env_length)) / 2
            down = (dist < torch.norm(self.commands[env_ids, :2], dim=1) * _f(np.float32(P.max_episode_length_s)) * 0.5) * ~up
            self.terrain_levels[env_ids] += 1 * up - 1 * down
            ctx = rng.DrawContext(P.seed, step, rng.CURRICULUM, ids_np, P.env_id_offset)
            rnd = torch.from_numpy(ctx.randint((len(env_ids),), P.max_terrain_level))
            self.terrain_levels[env_ids] = torch.where(self.terrain_levels[env_ids] >= P.max_terrain_level, rnd,
                                                       torch.clip(self.terrain_levels[env_ids], 0))
            self.env_origins[env_ids] = self.terrain_origins[self.terrain_levels[env_ids], self.terrain_types[env_ids]]
        ctx = rng.DrawContext(P.seed, step, rng.RESET_DOF, ids_np, P.env_id_offset)
        u = torch.from_numpy(ctx.uniform((len(env_ids), self.D)))
        self.dof_pos[env_ids] = self.default_dof_pos * ((1.5 - 0.5) * u + 0.5)
        self.dof_vel[env_ids] = 0.
        ctx = rng.DrawContext(P.seed, step, rng.RESET_ROOT, ids_np, P.env_id_offset, 0 if P.custom_origins else 2)
        self.root_states[env_ids] = self.base_init_state
        self.root_states[env_ids, :3] += self.env_origins[env_ids]
        if P.custom_origins:
            self.root_states[env_ids, :2] += (1. - (-1.)) * torch.from_numpy(ctx.uniform((len(env_ids), 2))) + (-1.)
        self.root_states[env_ids, 7:13] = (0.5 - (-0.5)) * torch.from_numpy(ctx.uniform((len(env_ids), 6))) + (-0.5)
        for buf in (self.last_actions, self.last_dof_vel, self.feet_air_time, self.last_feet_air_time,
                    self.current_max_feet_height, self.last_max_feet_height):
            buf[env_ids] = 0.
        self.episode_length_buf[env_ids] = 0
        self.reset_buf[env_ids] = 1
        R = self.episode_sums.shape[0]
        for k in range(R):
            self.episode_stats[k] = torch.mean(self.episode_sums[k][env_ids]) / _f(np.float32(P.max_episode_length_s))
            self.episode_sums[k][env_ids] = 0.
        if P.terrain_curriculum:
            self.episode_stats[R] = torch.mean(self.terrain_levels.float())
        self.episode_stats[R + 1] = float(len(env_ids))
        if P.send_timeouts:     # PF:287-288: the dict entry is (re)bound only here, i.e. on steps with >= 1 reset
            self.extras_time_outs = self.time_out_buf

    # -------------------------------------------------------------------------------- PF:320-382
    def compute_observations(self):
        P = self.P
        g = lambda x: _f(np.float32(x))
        base = torch.cat((self.base_ang_vel * g(P.obs_scale_ang_vel), self.projected_gravity,
                          (self.dof_pos - self.default_dof_pos) * g(P.obs_scale_dof_pos),
                          self.dof_vel * g(P.obs_scale_dof_vel), self.actions,
                          self.commands[:, :3] * self.commands_scale), dim=-1)
        heights = None
        if P.measure_heights_actor or P.measure_heights_critic:
            heights = torch.clip(self.root_states[:, 2].unsqueeze(1) - 0.5 - self.measured_heights, -1, 1.) * g(P.obs_scale_height)
        self.obs_buf = torch.cat((base, heights), dim=-1) if P.measure_heights_actor else base
        if P.num_priv_obs:
            self.priv_obs_buf = torch.cat((base, heights), dim=-1) if P.measure_heights_critic else base.clone()
        if P.add_noise:
            ids = np.arange(self.N)
            ctx = rng.DrawContext(P.seed, self.common_step_counter, rng.OBS_NOISE, ids, P.env_id_offset)
            n_obs = self.obs_buf.shape[1]
            noise = (2 * torch.from_numpy(ctx.uniform(self.obs_buf.shape)) - 1) * self.noise_vec[:n_obs]
            self.obs_buf += noise
            if P.num_priv_obs:
                extra = (2 * torch.from_numpy(ctx.uniform((self.N, P.num_priv_obs - n_obs))) - 1) * self.noise_vec[n_obs:]
                self.priv_obs_buf += torch.cat((noise, extra), dim=1)


def heights_native_eager(orc, device):
    """``_get_heights`` (PF:1048-1075, 1100-1112) of a RefEnv with the tensors on ``device`` and the division of PF:1102
    left to torch's own ``/`` operator there: ATen's CPU kernel divides, its CUDA kernel multiplies by the reciprocal of
    the scalar.  Used by the GPU tests to pin PF_DIV_RECIPROCAL to what the reference does on a CUDA device."""
    P = orc.P
    root = orc.root_states.to(device)
    pts0 = orc.height_points.to(device)
    hs_t = orc.height_samples.to(device)
    quat = root[:, 3:7].repeat(1, orc.num_height_points)
    qy = quat.clone().view(-1, 4)
    qy[:, :2] = 0.
    qy = normalize(qy)
    points = quat_apply(qy, pts0) + (root[:, :3]).unsqueeze(1)
    points = points + _f(P.border_size)
    points = (points / _f(np.float32(P.horizontal_scale))).long()
    px = torch.clip(points[:, :, 0].view(-1), 0, hs_t.shape[0] - 2)
    py = torch.clip(points[:, :, 1].view(-1), 0, hs_t.shape[1] - 2)
    h = torch.min(hs_t[px, py], hs_t[px + 1, py])
    h = torch.min(h, hs_t[px, py + 1])
    return h.view(orc.N, -1) * _f(np.float32(P.vertical_scale))
