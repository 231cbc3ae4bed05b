"""``PointFoot`` -- drop-in for legged_gym/envs/pointfoot/point_foot.py::PointFoot on the per-step path.

Same constructor signature, attributes and VecEnv protocol (``step(actions) -> (obs, privileged_obs, rew, reset,
extras)``, ``reset()``, ``get_observations()``, ``get_privileged_observations()``; point_foot.py:17, 131-142,
170-193), but every per-step computation is a launch into libpointfoot_b200.so (include/pointfoot_b200.h): there
is no torch arithmetic on the path and no CPU fallback.  Physics stays behind the ``gym`` object exactly as in the
reference (point_foot.py:182-185, 200-202): pass a real ``isaacgym.gymapi`` Gym, or leave the default
``SyntheticGym`` which hands out seeded state tensors (Isaac Gym is a closed binary that cannot be installed here).

Differences from the reference that are deliberate:
  * the device is taken from ``sim_device`` (the fork hard-codes 'cpu', point_foot.py:41,50);
  * random draws follow the counter-based contract of csrc/pf_rng.cuh instead of torch's global Philox stream;
  * ``torch.save`` of the noise buffer every step (point_foot.py:334) is not replicated.
"""
import ctypes as C

import numpy as np
import torch

from . import capi, params as pf_params, synth


class SyntheticGym:
    """The ``gym.*`` calls the hot path makes (point_foot.py:182-185, 200-202, 562-564, 586-588, 606-611), inert,
    over synthetic state tensors.  ``feeder(gym, substep)`` may rewrite the state tensors in place on
    ``simulate`` to stand in for PhysX."""
    needs_host_reset_count = False      # a real Gym needs len(env_ids) on the host for the indexed setters

    def __init__(self, state, feeder=None):
        self.state = state
        self.feeder = feeder
        self.substep = 0

    def acquire_actor_root_state_tensor(self, sim): return self.state["root_states"]
    def acquire_dof_state_tensor(self, sim): return self.state["dof_state"]
    def acquire_net_contact_force_tensor(self, sim): return self.state["contact_forces"]
    def acquire_rigid_body_state_tensor(self, sim): return self.state["rigid_body_states"]

    def simulate(self, sim):
        if self.feeder is not None:
            self.feeder(self, self.substep)
        self.substep += 1

    def _inert(self, *a, **k): return True
    fetch_results = refresh_dof_state_tensor = refresh_actor_root_state_tensor = _inert
    refresh_net_contact_force_tensor = refresh_rigid_body_state_tensor = _inert
    set_dof_actuation_force_tensor = set_dof_state_tensor_indexed = set_actor_root_state_tensor_indexed = _inert
    apply_rigid_body_force_tensors = _inert


class _Wrap:
    """gymtorch.wrap_tensor / unwrap_tensor stand-ins (identity) for the synthetic backend."""
    @staticmethod
    def wrap_tensor(t): return t
    @staticmethod
    def unwrap_tensor(t): return t


class PointFoot:
    def __init__(self, cfg, sim_params=None, physics_engine=None, sim_device="cuda:0", headless=True,
                 gym=None, gymtorch=None, state=None, seed=1, env_id_offset=0, resample_commands=False,
                 keep_measured_heights=False):
        self.cfg = cfg
        self._keep_measured_heights = bool(keep_measured_heights)
        self.sim_params = sim_params
        self.physics_engine = physics_engine
        self.headless = headless
        self.viewer = None                      # graphics API: unavailable on this box (headless only)
        self.enable_viewer_sync = True
        self.debug_viz = False
        self.init_done = False
        self.device = torch.device(sim_device)
        self.sim_device = str(sim_device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("pointfoot_b200.PointFoot needs a CUDA device (no CPU fallback); got " + str(sim_device))
        self._lib = capi.load()
        # multi-GPU block sharding: global id of local env 0 (the RNG contract keys on global ids; the runner hands the
        # same offset to the policy's sampling kernel) and the contract's seed
        self.env_id_offset, self.seed = int(env_id_offset), int(seed)
        N = self.num_envs = cfg.env.num_envs
        self.num_obs = cfg.env.num_propriceptive_obs
        self.num_privileged_obs = cfg.env.num_privileged_obs
        self.num_actions = cfg.env.num_actions
        task = "pointfoot_rough" if cfg.terrain.mesh_type in ("heightfield", "trimesh") else "pointfoot_flat"

        # ---- physics backend (create_sim, point_foot.py:384-401: closed PhysX, or the synthetic stand-in) ----
        if gym is None:
            if state is None:
                state = synth.make_state(N, task, seed=seed, env_id_offset=env_id_offset)
            state = {k: (v.to(self.device).contiguous() if torch.is_tensor(v) else v) for k, v in state.items()}
            gym, gymtorch = SyntheticGym(state), _Wrap
        self.gym, self._gymtorch, self.sim = gym, gymtorch or _Wrap, getattr(gym, "sim", None)
        self._state = state or {}
        self.envs, self.actor_handles = [None] * N, [None] * N

        # ---- constants: _parse_cfg / _process_dof_props / _prepare_reward_function, flattened ----
        hf = self._state.get("height_samples")
        self.height_samples = hf
        P, noise, pts, names = pf_params.build_params(
            cfg, N, seed=seed, env_id_offset=env_id_offset, hf_shape=tuple(hf.shape) if hf is not None else (0, 0),
            resample_commands=resample_commands)
        self._P, self.reward_names = P, names
        self.noise_scale_vec = noise.to(self.device)
        self.dt = cfg.control.decimation * cfg.sim.dt
        self.obs_scales = cfg.normalization.obs_scales
        self.reward_scales = {n: float(P.reward_scales[i]) for i, n in enumerate(names[:P.num_reward_terms])}
        self.command_ranges = pf_params.class_to_dict(cfg.commands.ranges)
        self.max_episode_length_s = cfg.env.episode_length_s
        self.max_episode_length = np.ceil(self.max_episode_length_s / self.dt)
        self.push_interval = int(np.ceil(cfg.domain_rand.push_interval_s / self.dt))
        self.num_dof = self.num_dofs = P.num_dof
        self.num_bodies = P.num_bodies
        self.dof_names = list(pf_params.RobotModel.dof_names)
        self.custom_origins = bool(P.custom_origins)
        self._include_feet_height_rewards = bool(P.include_feet_height_rewards)
        self._init_buffers(names)

        # what pf_post_physics_step launches besides the step kernel (bench.py names them in its roofline block)
        self.tail_kernel_name = " + pf_finalize_kernel"
        self._ctx = C.c_void_p()
        pts = pts.contiguous()
        capi.check(self._lib.pf_ctx_create(C.byref(P), C.c_void_p(noise.contiguous().data_ptr()),
                                           C.c_void_p(pts.data_ptr() if pts.numel() else None), C.byref(self._ctx)),
                   "pf_ctx_create")
        self.init_done = True

    # ------------------------------------------------------------------ buffers (point_foot.py:56-77, 698-794)
    def _init_buffers(self, names):
        P, N, dev = self._P, self.num_envs, self.device
        D, NB = P.num_dof, P.num_bodies
        z = lambda *s, dt=torch.float: torch.zeros(*s, dtype=dt, device=dev)
        wrap = self._gymtorch.wrap_tensor
        self.root_states = wrap(self.gym.acquire_actor_root_state_tensor(self.sim))
        self.dof_state = wrap(self.gym.acquire_dof_state_tensor(self.sim))
        self.dof_pos = self.dof_state.view(N, D, 2)[..., 0]
        self.dof_vel = self.dof_state.view(N, D, 2)[..., 1]
        self.base_quat = self.root_states[:, 3:7]
        self.rigid_body_states = wrap(self.gym.acquire_rigid_body_state_tensor(self.sim)).view(N, NB, -1)
        self.contact_forces = wrap(self.gym.acquire_net_contact_force_tensor(self.sim)).view(N, -1, 3)
        self.feet_indices = torch.tensor([P.feet_indices[i] for i in range(2)], device=dev)
        self.termination_contact_indices = torch.tensor([P.termination_indices[i] for i in range(P.num_termination)], device=dev)
        self.penalised_contact_indices = torch.tensor([P.penalised_indices[i] for i in range(P.num_penalised)], device=dev)
        self.common_step_counter = 0
        self._step_counter = torch.zeros(1, dtype=torch.int64, device=dev)
        self.extras = {}
        self.proprioceptive_obs_buf = z(N, self.num_obs)
        self.privileged_obs_buf = z(N, self.num_privileged_obs) if self.num_privileged_obs is not None else None
        self.rew_buf = z(N)
        self.reset_buf = torch.ones(N, dtype=torch.bool, device=dev)      # long->bool at point_foot.py:68 vs 242
        self.episode_length_buf = z(N, dt=torch.long)
        self.time_out_buf, self.terminate = z(N, dt=torch.bool), z(N, dt=torch.bool)
        self.torques, self.actions, self.last_actions = z(N, D), z(N, D), z(N, D)
        self.last_dof_vel, self.last_root_vel = z(N, D), z(N, 6)
        self.commands = z(N, P.num_commands)
        if "commands" in self._state:
            self.commands.copy_(self._state["commands"])
        self.commands_scale = torch.tensor([P.commands_scale[i] for i in range(3)], device=dev)
        self.feet_air_time, self.last_feet_air_time = z(N, 2), z(N, 2)
        self.contact_filt, self.first_contact, self.last_contacts = (z(N, 2, dt=torch.bool) for _ in range(3))
        self.feet_height, self.last_max_feet_height, self.current_max_feet_height = z(N, 2), z(N, 2), z(N, 2)
        self.rigid_body_external_forces, self.rigid_body_external_torques = z(N, NB, 3), z(N, NB, 3)
        self.base_lin_vel, self.base_ang_vel, self.projected_gravity = z(N, 3), z(N, 3), z(N, 3)
        self.default_dof_pos = torch.tensor([[P.default_dof_pos[i] for i in range(D)]], device=dev)
        self.p_gains = torch.tensor([P.p_gains[i] for i in range(D)], device=dev)
        self.d_gains = torch.tensor([P.d_gains[i] for i in range(D)], device=dev)
        self.torque_limits = torch.tensor([P.torque_limits[i] for i in range(D)], device=dev)
        self.base_mass = self._state.get("base_mass", torch.full((N,), pf_params.RobotModel.base_mass, device=dev))
        self._refresh_max_push_force()
        # PF:775 / 212-213 keep the scan as an attribute, but its only readers on the step path (height observation columns,
        # base-height reward) are fused into the kernels; the [N,P] fp32 copy (127 MB per step at 262,144 envs) is written
        # only for a caller that asks (keep_measured_heights=True: debug drawing, parity tests)
        self._measured_heights = z(N, P.num_height_points) if (P.num_height_points and self._keep_measured_heights) else None
        R = P.num_reward_terms + P.has_termination_term
        self._episode_sums = z(R, N)
        self.episode_sums = {n: self._episode_sums[i] for i, n in enumerate(names)}
        self._episode_stats = z(R + 2)
        self._reset_ids = z(N, dt=torch.int32)
        self._reset_count = z(1, dt=torch.int32)
        self._time_outs = z(N, dt=torch.bool)
        # terrain / origins (point_foot.py:970-997)
        if self.custom_origins:
            self.terrain_origins = self._state["terrain_origins"]
            self.terrain_levels = self._state["terrain_levels"].clone()
            self.terrain_types = self._state["terrain_types"]
            self.max_terrain_level = P.max_terrain_level
            self.env_origins = self.terrain_origins[self.terrain_levels, self.terrain_types].clone().contiguous()
        else:
            self.terrain_origins = self.terrain_levels = self.terrain_types = None
            o = z(N, 3)
            num_cols = np.floor(np.sqrt(N)); num_rows = np.ceil(N / num_cols)
            xx, yy = torch.meshgrid(torch.arange(num_rows), torch.arange(num_cols), indexing="ij")
            sp = self.cfg.env.env_spacing
            o[:, 0] = (sp * xx.flatten()[:N]).to(dev); o[:, 1] = (sp * yy.flatten()[:N]).to(dev)
            self.env_origins = o
        b = capi.PfEnvBuffers()
        fields = dict(
            root_states=self.root_states, dof_state=self.dof_state, contact_forces=self.contact_forces,
            rigid_body_states=self.rigid_body_states, actions=self.actions, torques=self.torques,
            commands=self.commands, episode_length_buf=self.episode_length_buf, base_lin_vel=self.base_lin_vel,
            base_ang_vel=self.base_ang_vel, projected_gravity=self.projected_gravity,
            last_actions=self.last_actions, last_dof_vel=self.last_dof_vel, last_root_vel=self.last_root_vel,
            feet_air_time=self.feet_air_time, last_feet_air_time=self.last_feet_air_time,
            current_max_feet_height=self.current_max_feet_height, last_max_feet_height=self.last_max_feet_height,
            feet_height=self.feet_height, contact_filt=self.contact_filt, first_contact=self.first_contact,
            last_contacts=self.last_contacts, episode_sums=self._episode_sums, rew_buf=self.rew_buf,
            reset_buf=self.reset_buf, time_out_buf=self.time_out_buf, terminate=self.terminate,
            obs_buf=self.proprioceptive_obs_buf, priv_obs_buf=self.privileged_obs_buf,
            measured_heights=self._measured_heights,
            height_samples=self.height_samples, terrain_origins=self.terrain_origins,
            terrain_levels=self.terrain_levels, terrain_types=self.terrain_types, env_origins=self.env_origins,
            reset_ids=self._reset_ids, reset_count=self._reset_count, episode_stats=self._episode_stats,
            extras_time_outs=self._time_outs, step_counter=self._step_counter)
        for name in capi.ENV_BUFFER_FIELDS:
            t = fields[name]
            if t is not None and not t.is_contiguous():
                raise RuntimeError(f"buffer {name} must be contiguous")
            setattr(b, name, None if t is None else t.data_ptr())
        self._bufs, self._buf_refs = b, fields
        # extras: 0-d views, refreshed in place by the finalize kernel only on steps with >= 1 reset (PF:259, 276-288)
        ep = {"rew_" + n: self._episode_stats[i] for i, n in enumerate(names)}
        if self._P.terrain_curriculum:
            ep["terrain_level"] = self._episode_stats[R]
        if self.cfg.commands.curriculum:
            ep["max_command_x"] = self.command_ranges["lin_vel_x"][1]
        self.extras["episode"] = ep
        if self.cfg.env.send_timeouts:
            self.extras["time_outs"] = self._time_outs

    @property
    def measured_heights(self):
        """PF:775 (the int 0 when the task measures nothing), PF:1054-1085"""
        if not self._P.num_height_points:
            return 0
        if self._measured_heights is None:
            raise AttributeError("measured_heights is not materialised on the step path by default; construct the env with "
                                 "keep_measured_heights=True (the height observations and the base-height reward do not need it)")
        return self._measured_heights

    # ------------------------------------------------------------------ VecEnv protocol
    def get_observations(self):
        return self.proprioceptive_obs_buf

    def get_privileged_observations(self):
        return self.privileged_obs_buf

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def reset(self):
        """point_foot.py:137-142: reset every env, then one zero-action step."""
        capi.check(self._lib.pf_reset_all(self._ctx, C.byref(self._bufs), C.c_int64(self.common_step_counter),
                                          self._stream()), "pf_reset_all")
        self._notify_gym_of_resets()
        obs, priv, _, _, _ = self.step(torch.zeros(self.num_envs, self.num_actions, device=self.device))
        return obs, priv

    def step(self, actions):
        """point_foot.py:170-193."""
        lib, ctx, b, gym, sim = self._lib, self._ctx, C.byref(self._bufs), self.gym, self.sim
        st = self._stream()
        a = actions.to(self.device, torch.float).contiguous()
        self.render()
        for i in range(self.cfg.control.decimation):
            capi.check(lib.pf_pd_torques(ctx, b, C.c_void_p(a.data_ptr()) if i == 0 else None, st), "pf_pd_torques")
            gym.set_dof_actuation_force_tensor(sim, self._gymtorch.unwrap_tensor(self.torques))
            gym.simulate(sim)
            gym.fetch_results(sim, True)
            gym.refresh_dof_state_tensor(sim)
        self.post_physics_step()
        return (self.proprioceptive_obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf, self.extras)

    def post_physics_step(self):
        """point_foot.py:195-230 (one fused launch + the finalize launch)."""
        gym, sim = self.gym, self.sim
        gym.refresh_actor_root_state_tensor(sim)
        gym.refresh_net_contact_force_tensor(sim)
        gym.refresh_rigid_body_state_tensor(sim)
        self.common_step_counter += 1
        st = self._stream()
        # (a step being captured into a CUDA graph never carries the push: GraphedStep runs push steps eagerly)
        if (self.cfg.domain_rand.push_robots and self.common_step_counter % self.push_interval == 0
                and not getattr(self, "_in_capture", False)):
            self._push_robots(st)
        if getattr(self, "_defer_tail", False):
            # GraphedStep with a policy: the step's tail (reset id list, count, extras["episode"]) runs on the context's
            # side stream beside the policy kernel; join_tail() makes this stream wait for it
            capi.check(self._lib.pf_post_physics_step_begin(self._ctx, C.byref(self._bufs), C.c_int64(self.common_step_counter), st),
                       "pf_post_physics_step_begin")
            return
        capi.check(self._lib.pf_post_physics_step(self._ctx, C.byref(self._bufs), C.c_int64(self.common_step_counter), st),
                   "pf_post_physics_step")
        self._notify_gym_of_resets()
        self._maybe_command_curriculum()

    def _maybe_command_curriculum(self):
        if self._command_curriculum_due() and int(self._reset_count.item()) > 0:   # reset_idx returns early without resets (PF:259)
            self.update_command_curriculum()

    def _command_curriculum_due(self):
        """point_foot.py:265: only on steps where common_step_counter % max_episode_length == 0 (every 1000 steps)"""
        return bool(self.cfg.commands.curriculum) and self.common_step_counter % self.max_episode_length == 0

    def join_tail(self):
        """make the current stream wait for a deferred step tail (pf_post_physics_step_begin / _join)"""
        capi.check(self._lib.pf_post_physics_step_join(self._ctx, self._stream()), "pf_post_physics_step_join")

    def _refresh_max_push_force(self):
        """point_foot.py:592-596 evaluates base_mass.mean().item() * max_push_vel_xy / dt on every push; base_mass only
        changes through update_added_mass_and_base_com, so the python double is kept (no host sync on the step path,
        and a push step can sit inside a stream capture)."""
        # (the mean is taken where the reference takes it: on the host, point_foot.py:41,50 -- ATen's CPU and CUDA
        # reductions round differently from a few thousand envs on)
        self._max_push_force = (float(self.base_mass.detach().cpu().mean().item()) * self.cfg.domain_rand.max_push_vel_xy
                                / self.cfg.sim.dt)

    def _push_robots(self, st):
        """point_foot.py:590-611."""
        m = self._max_push_force
        capi.check(self._lib.pf_push_forces(self._ctx, C.byref(self._bufs),
                                            C.c_void_p(self.rigid_body_external_forces.data_ptr()), C.c_float(m),
                                            C.c_int64(self.common_step_counter), st), "pf_push_forces")
        self.gym.apply_rigid_body_force_tensors(self.sim, self._gymtorch.unwrap_tensor(self.rigid_body_external_forces),
                                                self._gymtorch.unwrap_tensor(self.rigid_body_external_torques), 0)

    def _notify_gym_of_resets(self):
        """point_foot.py:561-564, 585-588: the indexed setters want an int32 id list and a HOST count; the rows
        themselves were already written in place by the kernel.  The synthetic backend needs neither."""
        if not getattr(self.gym, "needs_host_reset_count", True):
            return
        n = int(self._reset_count.item())
        if n == 0:
            return
        ids = self._gymtorch.unwrap_tensor(self._reset_ids)
        self.gym.set_dof_state_tensor_indexed(self.sim, self._gymtorch.unwrap_tensor(self.dof_state), ids, n)
        self.gym.set_actor_root_state_tensor_indexed(self.sim, self._gymtorch.unwrap_tensor(self.root_states), ids, n)

    # ------------------------------------------------------------------ the rest of the public surface
    def render(self, sync_frame_time=True):
        return None                              # viewer: unavailable on this box

    def update_cmd(self, cmd):
        """point_foot.py:123-129 (sys-id scripts drive env 0 by hand)."""
        self.commands[0, :3] = torch.as_tensor(cmd[:3], dtype=torch.float, device=self.device)

    def update_frictions(self, fric, env_handle, actor_handle):
        """point_foot.py:103-108: a pass-through to the physics backend (nothing on the per-step path reads it)."""
        dof_props = self.gym.get_actor_dof_properties(env_handle, actor_handle)
        for i in range(len(dof_props["friction"])):
            dof_props["friction"][i] = fric[i]
        self.gym.set_actor_dof_properties(env_handle, actor_handle, dof_props)

    def update_added_mass_and_base_com(self, added_mass, added_com, env_handle, actor_handle):
        """point_foot.py:110-121: pass-through to the physics backend; ``base_mass`` becomes the 0-d tensor the
        reference leaves behind (so the push force scale follows, point_foot.py:592-596)."""
        body_props = self.gym.get_actor_rigid_body_properties(env_handle, actor_handle)
        body_props[0].mass += added_mass
        self.base_mass = torch.tensor(body_props[0].mass, dtype=torch.float)
        body_props[0].com.x += added_com[0]
        body_props[0].com.y += added_com[1]
        body_props[0].com.z += added_com[2]
        self.gym.set_actor_rigid_body_properties(env_handle, actor_handle, body_props, recomputeInertia=True)
        self._refresh_max_push_force()

    def update_command_curriculum(self, env_ids=None):
        """point_foot.py:638-650: widen the lin_vel_x command range when the mean tracking reward of the envs that just
        reset exceeds 80 % of its maximum.  ``extras["episode"]["rew_tracking_lin_vel"]`` already holds that mean
        (divided by max_episode_length_s, point_foot.py:279-280).  The new range goes into the context (the LR
        command resampling of LR:337-351 draws from it) and into ``command_ranges``."""
        if "tracking_lin_vel" not in self.reward_names:
            raise KeyError("tracking_lin_vel")           # the reference indexes episode_sums / reward_scales by that name
        k = self.reward_names.index("tracking_lin_vel")
        mean = float(self._episode_stats[k].item()) * self.max_episode_length_s
        if mean / self.max_episode_length > 0.8 * self.reward_scales["tracking_lin_vel"]:
            r, mx = self.command_ranges["lin_vel_x"], self.cfg.commands.max_curriculum
            r[0] = float(np.clip(r[0] - 0.5, -mx, 0.))
            r[1] = float(np.clip(r[1] + 0.5, 0., mx))
            self._push_command_ranges()

    def _push_command_ranges(self):
        cr = self.command_ranges
        arr = (C.c_float * 8)(*[float(v) for key in ("lin_vel_x", "lin_vel_y", "ang_vel_yaw", "heading") for v in cr[key]])
        capi.check(self._lib.pf_set_command_ranges(self._ctx, arr), "pf_set_command_ranges")
        for i, key in enumerate(("lin_vel_x", "lin_vel_y", "ang_vel_yaw", "heading")):
            self._P.cmd_range[i][0], self._P.cmd_range[i][1] = cr[key]
        self.graph_epoch = getattr(self, "graph_epoch", 0) + 1      # constants changed: a captured step must be re-captured

    @property
    def reset_env_ids(self):
        """ascending ids of the envs that reset in the last step (device int32 view; syncs for the count)"""
        return self._reset_ids[:int(self._reset_count.item())]

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.pf_ctx_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GraphedStep:
    """``env.step(actions)`` (+ an optional ``extra()`` such as the policy forward that produces the next actions)
    captured once into a CUDA graph and replayed: the per-step cost becomes one graph launch instead of six or more
    kernel launches.  Only valid when nothing on the host has to happen inside the step, i.e. for backends that
    do not need the reset count on the host (SyntheticGym, or a GPU-pipeline gym with device-side setters).
    Steps on which a push is due (point_foot.py:492-493) run eagerly."""

    def __init__(self, env, actions, extra=None):
        if getattr(env.gym, "needs_host_reset_count", True):
            raise RuntimeError("GraphedStep needs a backend without host work inside the step")
        self.env, self.actions, self.extra = env, actions, extra
        side = torch.cuda.Stream(device=env.device)
        side.wait_stream(torch.cuda.current_stream(env.device))
        with torch.cuda.stream(side):
            self._eager()                      # warm-up on the side stream (lazy module loads, allocator)
        torch.cuda.current_stream(env.device).wait_stream(side)
        torch.cuda.synchronize(env.device)
        self._capture()

    def _capture(self):
        env = self.env
        self.graph = torch.cuda.CUDAGraph()
        host_counter = env.common_step_counter
        self._capturing = env._in_capture = True
        n0 = env._lib.pf_debug_launch_count()
        try:
            with torch.cuda.graph(self.graph):
                self._eager()
        finally:
            self._capturing = env._in_capture = False
        self.launches_per_step = int(env._lib.pf_debug_launch_count() - n0)     # kernels of this library per replay
        env.common_step_counter = host_counter     # capture does not execute
        self._epoch = getattr(env, "graph_epoch", 0)

    def _eager(self):
        if self.extra is None:
            return self.env.step(self.actions)
        # step + policy: the policy's act kernel needs the observations only, so the step's tail (reset ids, count,
        # episode statistics) is left running beside it and joined before the graph ends
        self.env._defer_tail = True
        try:
            out = self.env.step(self.actions)
            self.extra()
            self.env.join_tail()
        finally:
            self.env._defer_tail = False
        if not getattr(self, "_capturing", False):
            self.env._maybe_command_curriculum()
        return out

    def _push_due(self):
        """steps with host-side work (a push, point_foot.py:492-493; the command curriculum, :265) run eagerly"""
        env = self.env
        nxt = env.common_step_counter + 1
        return ((env.cfg.domain_rand.push_robots and nxt % env.push_interval == 0)
                or (bool(env.cfg.commands.curriculum) and nxt % env.max_episode_length == 0))

    def __call__(self):
        env = self.env
        if self._push_due():
            self._eager()
            if getattr(env, "graph_epoch", 0) != self._epoch:      # kernel constants changed (command ranges): re-capture
                torch.cuda.synchronize(env.device)
                self._capture()
        else:
            self.graph.replay()
            env.common_step_counter += 1
        return (env.proprioceptive_obs_buf, env.privileged_obs_buf, env.rew_buf, env.reset_buf, env.extras)
