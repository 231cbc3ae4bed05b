"""Run the UNMODIFIED reference ``point_foot.py`` behind ``sys.modules`` stubs.  ORACLE / TEST CODE.

Only usable where /root/reference exists (the build container).  It pins ``oracle.pointfoot_ref``
and generates ``tests/golden/*.npz`` (tests/golden/make_golden.py).  Nothing on the GPU box imports it.

What is stubbed: ``isaacgym{,.gymapi,.gymutil,.gymtorch,.terrain_utils}`` = inert mocks,
``isaacgym.torch_utils`` = oracle.isaacgym_torch_utils, ``rsl_rl*`` and ``matplotlib*`` = empty modules.
``PointFoot`` is built with ``object.__new__`` and the fields its ctor / ``_create_envs`` would have filled
from PhysX (point_foot.py:56-77, 900-968) are set from the URDF constants instead; then the reference's own
``_parse_cfg`` / ``_init_buffers`` / ``_prepare_reward_function`` run.  Random draws are routed to the
counter-based contract (oracle.rng) by wrapping the five methods that draw; ``torch.save``
(point_foot.py:334) is neutralised.
"""
import os
import sys
import types
import contextlib

import numpy as np
import torch

from . import rng
from . import isaacgym_torch_utils as _tu

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _default_root():
    """$POINTFOOT_REFERENCE_ROOT, else the reference tree of the build container, else the offline install that travels
    with the repo snapshot (baseline/_ref, made by baseline/install_reference.py)."""
    env = os.environ.get("POINTFOOT_REFERENCE_ROOT")
    if env:
        return env
    for root in ("/root/reference", os.path.join(_REPO, "baseline", "_ref")):
        if os.path.isfile(os.path.join(root, "legged_gym/envs/pointfoot/point_foot.py")):
            return root
    return "/root/reference"


REFERENCE_ROOT = _default_root()


def reference_available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "legged_gym/envs/pointfoot/point_foot.py"))


class _Anything:
    """Inert object: any attribute / call returns another inert object."""
    def __init__(self, *a, **k):
        pass

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything()

    def __call__(self, *a, **k):
        return _Anything()

    def __iter__(self):
        return iter(())


class FakeGym:
    """The 12 ``gym.*`` calls the hot path makes (point_foot.py:182-185, 200-202, 562-564, 586-588, 606-611)
    are inert; the acquire_* calls hand out the synthetic state tensors."""
    def __init__(self, tensors):
        self.t = tensors
        self.calls = []

    def acquire_actor_root_state_tensor(self, sim): return self.t["root_states"]
    def acquire_dof_state_tensor(self, sim): return self.t["dof_state"]
    def acquire_net_contact_force_tensor(self, sim): return self.t["contact_forces"]
    def acquire_rigid_body_state_tensor(self, sim): return self.t["rigid_body_states"]

    def __getattr__(self, name):
        def _inert(*a, **k):
            self.calls.append(name)
            return None
        return _inert


_installed = False


def install_stubs():
    global _installed
    if _installed:
        return
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")

    import importlib.machinery

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__spec__ = importlib.machinery.ModuleSpec(name, None)     # find_spec() on a stub must not raise
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class _ModAny(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _Anything

    def anymod(name):
        m = _ModAny(name)
        m.__spec__ = importlib.machinery.ModuleSpec(name, None)
        sys.modules[name] = m
        return m

    ig = mod("isaacgym")
    ig.gymapi = anymod("isaacgym.gymapi")
    ig.gymutil = anymod("isaacgym.gymutil")
    ig.terrain_utils = anymod("isaacgym.terrain_utils")
    ig.gymtorch = mod("isaacgym.gymtorch", wrap_tensor=lambda t: t, unwrap_tensor=lambda t: t)
    tu = mod("isaacgym.torch_utils")
    for k in ("to_torch", "get_axis_params", "normalize", "quat_apply", "quat_rotate",
              "quat_rotate_inverse", "torch_rand_float"):
        setattr(tu, k, getattr(_tu, k))
    tu.torch = torch
    tu.np = np            # point_foot.py uses ``np`` via ``from isaacgym.torch_utils import *`` (:470, :991)
    ig.torch_utils = tu
    for name in ("rsl_rl", "rsl_rl.env", "rsl_rl.runners", "rsl_rl.modules", "rsl_rl.algorithms",
                 "matplotlib", "matplotlib.pyplot", "onnxruntime"):
        anymod(name)
    sys.modules["rsl_rl.env"].VecEnv = object
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _installed = True


# --------------------------------------------------------------------------------------------
# PF_P441A constants PhysX would have reported (resources/robots/PF_P441A/urdf/PF_P441A.urdf:148,191,234,
# 390,433,476).  Body order after fixed-joint collapse: base, abad_L, hip_L, knee_L, foot_L, abad_R, hip_R,
# knee_R, foot_R.
DOF_NAMES = ["abad_L_Joint", "hip_L_Joint", "knee_L_Joint", "abad_R_Joint", "hip_R_Joint", "knee_R_Joint"]
BODY_NAMES = ["base_Link", "abad_L_Link", "hip_L_Link", "knee_L_Link", "foot_L_Link",
              "abad_R_Link", "hip_R_Link", "knee_R_Link", "foot_R_Link"]
DOF_LOWER = [-0.38397, -1.012291, -0.872665, -1.39626, -1.518436, -1.396263]
DOF_UPPER = [1.39626, 1.518436, 1.396263, 0.38397, 1.012291, 0.872665]
DOF_VELOCITY = [20.0, 20.0, 20.0, 20.0, 20.0, 50.0]
DOF_EFFORT = [80.0] * 6


@contextlib.contextmanager
def _patched(obj, name, value):
    old = getattr(obj, name)
    setattr(obj, name, value)
    try:
        yield
    finally:
        setattr(obj, name, old)


def _make_plain_class():
    """The reference PointFoot with torch's own generators (``torch.rand_like`` ...): what bench.py TIMES (on the host
    cores and, as the torch-eager bar, on the CUDA device).  The one change: the ``torch.save`` of the noise tensor that
    point_foot.py:334 does on every step (a leftover debug dump) is skipped -- timing disk writes
    would only flatter the comparison."""
    install_stubs()
    from legged_gym.envs.pointfoot.point_foot import PointFoot

    class TimedPointFoot(PointFoot):
        def _add_noise_to_obs(self):
            with _patched(torch, "save", lambda *a, **k: None):
                super()._add_noise_to_obs()

    return TimedPointFoot


def _make_class():
    """Subclass of the reference PointFoot whose only additions route RNG draws to the contract."""
    install_stubs()
    from legged_gym.envs.pointfoot.point_foot import PointFoot

    class RefPointFoot(PointFoot):
        rng_seed = 1
        env_id_offset = 0

        def _ctx(self, purpose, env_ids, lane0=0):
            ids = env_ids.cpu().numpy() if torch.is_tensor(env_ids) else np.asarray(env_ids)
            return rng.draw(rng.DrawContext(self.rng_seed, int(self.common_step_counter), purpose, ids,
                                            self.env_id_offset, lane0))

        def _add_noise_to_obs(self):
            def rand_like(t):
                return rng.active_uniform(t.shape, t.device)
            with self._ctx(rng.OBS_NOISE, np.arange(self.num_envs)), \
                    _patched(torch, "rand_like", rand_like), _patched(torch, "save", lambda *a, **k: None):
                super()._add_noise_to_obs()

        def _reset_dofs(self, env_ids):
            with self._ctx(rng.RESET_DOF, env_ids):
                super()._reset_dofs(env_ids)

        def _reset_root_states(self, env_ids):
            with self._ctx(rng.RESET_ROOT, env_ids, 0 if self.custom_origins else 2):
                super()._reset_root_states(env_ids)

        def _push_robots(self):
            with self._ctx(rng.PUSH, np.arange(self.num_envs)):
                super()._push_robots()

        def _update_terrain_curriculum(self, env_ids):
            def randint_like(t, high):
                return rng.active_randint(t.shape, high, t.device).to(t.dtype)
            with self._ctx(rng.CURRICULUM, env_ids), _patched(torch, "randint_like", randint_like):
                super()._update_terrain_curriculum(env_ids)

    return RefPointFoot


def make_cfg(task, num_envs, mutate=None):
    install_stubs()
    from legged_gym.envs.pointfoot.mixed_terrain.pointfoot_rough_config import PointFootRoughCfg
    from legged_gym.envs.pointfoot.flat.pointfoot_flat_config import PointFootFlatCfg
    cfg = {"pointfoot_rough": PointFootRoughCfg, "pointfoot_flat": PointFootFlatCfg}[task]()
    cfg.env.num_envs = num_envs
    if mutate is not None:
        mutate(cfg)
    return cfg


def build_reference_env(task, state, seed=1, mutate_cfg=None, env_id_offset=0, device="cpu", contract_rng=True):
    """Build the reference ``PointFoot`` over the synthetic tensors in ``state`` (a dict of CPU torch tensors,
    see pointfoot_b200.synth.make_state).  The four PhysX-owned tensors are used in place (the reference
    writes reset rows into them, point_foot.py:557-588).  ``device``: where the reference's tensors live ('cpu' is what
    the fork hard-codes, point_foot.py:41,50; 'cuda' is upstream legged_gym's default pipeline).  ``contract_rng=False``
    leaves torch's own generators in place (timing runs)."""
    cls = _make_class() if contract_rng else _make_plain_class()
    from types import SimpleNamespace
    if str(device) != "cpu":
        assert not contract_rng, "the counter-based RNG contract is a CPU (numpy) statement"
        state = {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in state.items()}
    N = int(state["root_states"].shape[0])
    cfg = make_cfg(task, N, mutate_cfg)
    env = object.__new__(cls)
    env.rng_seed, env.env_id_offset = seed, env_id_offset

    # ---- what __init__ sets (point_foot.py:30-77) ----
    env.terminate = None
    env.cfg = cfg
    env.sim_params = SimpleNamespace(dt=cfg.sim.dt, use_gpu_pipeline=False)
    env.height_samples = None
    env.debug_viz = False
    env.init_done = False
    env._parse_cfg()
    env.gym = FakeGym(state)
    env.sim = None
    env.physics_engine = None
    env.sim_device = env.device = str(device)
    env.headless = True
    env.num_envs = N
    env.num_obs = cfg.env.num_propriceptive_obs
    env.num_privileged_obs = cfg.env.num_privileged_obs
    env.num_actions = cfg.env.num_actions
    env.proprioceptive_obs_buf = torch.zeros(N, env.num_obs, device=device)
    env.rew_buf = torch.zeros(N, device=device)
    env.reset_buf = torch.ones(N, dtype=torch.long, device=device)
    env.episode_length_buf = torch.zeros(N, dtype=torch.long, device=device)
    env.time_out_buf = torch.zeros(N, dtype=torch.bool, device=device)
    env.privileged_obs_buf = (torch.zeros(N, env.num_privileged_obs, device=device)
                              if env.num_privileged_obs is not None else None)
    env.extras = {}
    env.enable_viewer_sync = True
    env.viewer = None

    # ---- what create_sim / _create_envs set from PhysX (point_foot.py:384-401, 870-968) ----
    env.up_axis_idx = 2
    env.num_dof = env.num_dofs = 6
    env.num_bodies = 9
    env.dof_names = list(DOF_NAMES)
    props = np.zeros(6, dtype=[("lower", "f4"), ("upper", "f4"), ("velocity", "f4"), ("effort", "f4")])
    props["lower"], props["upper"] = DOF_LOWER, DOF_UPPER
    props["velocity"], props["effort"] = DOF_VELOCITY, DOF_EFFORT
    env._process_dof_props(props, 0)                       # reference code: soft limits, point_foot.py:449-463

    def idx(names):
        out = []
        for n in names:
            out.extend(i for i, s in enumerate(BODY_NAMES) if n in s)
        return torch.tensor(out, dtype=torch.long, device=device)
    env.feet_indices = torch.tensor([i for i, s in enumerate(BODY_NAMES) if cfg.asset.foot_name in s], device=device)
    env.penalised_contact_indices = idx(cfg.asset.penalize_contacts_on)
    env.termination_contact_indices = idx(cfg.asset.terminate_after_contacts_on)
    isl = cfg.init_state
    env.base_init_state = _tu.to_torch(isl.pos + isl.rot + isl.lin_vel + isl.ang_vel, device=device)
    env.base_mass = state["base_mass"].clone()
    if cfg.terrain.mesh_type in ("heightfield", "trimesh"):
        env.terrain = SimpleNamespace(cfg=cfg.terrain, env_length=cfg.terrain.terrain_length,
                                      env_origins=state["terrain_origins"].cpu().numpy())
        env.height_samples = state["height_samples"]
    env._get_env_origins()                                   # reference code, point_foot.py:970-997
    if env.custom_origins:
        env.terrain_levels = state["terrain_levels"].clone()
        env.terrain_types = state["terrain_types"].clone()
        env.env_origins[:] = env.terrain_origins[env.terrain_levels, env.terrain_types]
    env.envs, env.actor_handles = [None] * N, [None] * N

    # ---- the reference's own buffer setup ----
    env._include_feet_height_rewards = env._check_if_include_feet_height_rewards()
    env._init_buffers()
    env._prepare_reward_function()
    env.init_done = True

    # ---- internal state the synthetic generator randomises ----
    env.commands[:] = state["commands"]
    env.episode_length_buf[:] = state["episode_length_buf"]
    for k in ("last_actions", "last_dof_vel", "feet_air_time", "last_feet_air_time", "contact_filt",
              "first_contact", "last_contacts", "current_max_feet_height", "last_max_feet_height"):
        if k in state:
            getattr(env, k)[:] = state[k]
    if "episode_sums" in state:      # rows in the C-ABI's order: summation order, 'termination' last (include/pointfoot_b200.h)
        keys = [k for k in env.episode_sums.keys() if k != "termination"]
        keys += ["termination"] if "termination" in env.episode_sums else []
        for i, name in enumerate(keys):
            env.episode_sums[name][:] = state["episode_sums"][i]
    env.common_step_counter = int(state.get("common_step_counter", 0))
    return env


# --------------------------------------------------------------------------------------------
def build_reference_anymal_torques(num_envs, device="cpu"):
    """The UNMODIFIED ``Anymal._compute_torques`` (legged_gym/envs/anymal_c/anymal.py:62-78) with the reference's own
    actuator network (resources/actuator_nets/anydrive_v3_lstm.pt), on an ``Anymal`` built with ``object.__new__``: only the
    attributes that method and ``Anymal._init_buffers``' own lines (anymal.py:43-50) touch are set.  Needs the reference
    SOURCE tree (the .pt file is a resource, not part of the installed package)."""
    install_stubs()
    from types import SimpleNamespace
    from legged_gym.envs.anymal_c.anymal import Anymal
    from legged_gym.envs.anymal_c.mixed_terrains.anymal_c_rough_config import AnymalCRoughCfg
    cfg = AnymalCRoughCfg()
    root = "/root/reference"
    env = object.__new__(Anymal)
    env.cfg, env.device, env.num_envs, env.num_actions = cfg, device, num_envs, 12
    env.actuator_network = torch.jit.load(cfg.control.actuator_net_file.format(LEGGED_GYM_ROOT_DIR=root)).to(device)
    env.default_dof_pos = torch.tensor([list(cfg.init_state.default_joint_angles.values())], dtype=torch.float, device=device)
    env.dof_state = torch.zeros(num_envs * 12, 2, device=device)
    env.dof_pos = env.dof_state.view(num_envs, 12, 2)[..., 0]
    env.dof_vel = env.dof_state.view(num_envs, 12, 2)[..., 1]
    # anymal.py:43-50, verbatim attribute set-up (super()._init_buffers() needs PhysX)
    env.sea_input = torch.zeros(num_envs * 12, 1, 2, device=device, requires_grad=False)
    env.sea_hidden_state = torch.zeros(2, num_envs * 12, 8, device=device, requires_grad=False)
    env.sea_cell_state = torch.zeros(2, num_envs * 12, 8, device=device, requires_grad=False)
    env.sea_hidden_state_per_env = env.sea_hidden_state.view(2, num_envs, 12, 8)
    env.sea_cell_state_per_env = env.sea_cell_state.view(2, num_envs, 12, 8)
    return env


# --------------------------------------------------------------------------------------------
# LeggedRobot base class with the a1 asset: constants PhysX would have reported for
# resources/robots/a1/urdf/a1.urdf after fixed-joint collapse (feet kept by dont_collapse, a1.urdf:187,309,431,553).
A1_DOF_NAMES = [f"{leg}_{j}_joint" for leg in ("FR", "FL", "RR", "RL") for j in ("hip", "thigh", "calf")]
A1_BODY_NAMES = ["base"] + [f"{leg}_{b}" for leg in ("FR", "FL", "RR", "RL") for b in ("hip", "thigh", "calf", "foot")]
A1_DOF_LOWER = [-0.802851455917, -1.0471975512, -2.69653369433] * 4
A1_DOF_UPPER = [0.802851455917, 4.18879020479, -0.916297857297] * 4
A1_DOF_VELOCITY = [52.4, 28.6, 28.6] * 4
A1_DOF_EFFORT = [20.0, 55.0, 55.0] * 4


def _make_lr_class():
    """Subclass of the reference LeggedRobot whose only additions route RNG draws to the contract (oracle.rng).
    ``_resample_commands`` is called from the callback (LR:326) and from reset_idx (LR:166) on the same step: two purposes."""
    install_stubs()
    from legged_gym.envs.base.legged_robot import LeggedRobot

    class RefLeggedRobot(LeggedRobot):
        rng_seed = 1
        env_id_offset = 0
        _in_reset = False

        def _ctx(self, purpose, env_ids, lane0=0):
            ids = env_ids.cpu().numpy() if torch.is_tensor(env_ids) else np.asarray(env_ids)
            return rng.draw(rng.DrawContext(self.rng_seed, int(self.common_step_counter), purpose, ids, self.env_id_offset, lane0))

        def compute_observations(self):
            def rand_like(t):
                return rng.active_uniform(t.shape, t.device)
            with self._ctx(rng.OBS_NOISE, np.arange(self.num_envs)), _patched(torch, "rand_like", rand_like):
                super().compute_observations()

        def reset_idx(self, env_ids):
            self._in_reset = True
            try:
                super().reset_idx(env_ids)
            finally:
                self._in_reset = False

        def _resample_commands(self, env_ids):
            if len(env_ids) == 0:            # (torch_rand_float of an empty shape draws nothing)
                return super()._resample_commands(env_ids)
            with self._ctx(rng.RESAMPLE_RESET if self._in_reset else rng.RESAMPLE_CMD, env_ids):
                super()._resample_commands(env_ids)

        def _reset_dofs(self, env_ids):
            with self._ctx(rng.RESET_DOF, env_ids):
                super()._reset_dofs(env_ids)

        def _reset_root_states(self, env_ids):
            with self._ctx(rng.RESET_ROOT, env_ids, 0 if self.custom_origins else 2):
                super()._reset_root_states(env_ids)

        def _push_robots(self):
            with self._ctx(rng.PUSH, np.arange(self.num_envs)):
                super()._push_robots()

        def _update_terrain_curriculum(self, env_ids):
            def randint_like(t, high):
                return rng.active_randint(t.shape, high, t.device).to(t.dtype)
            with self._ctx(rng.CURRICULUM, env_ids), _patched(torch, "randint_like", randint_like):
                super()._update_terrain_curriculum(env_ids)

    return RefLeggedRobot


def make_a1_cfg(num_envs, mutate=None):
    install_stubs()
    from legged_gym.envs.a1.a1_config import A1RoughCfg
    cfg = A1RoughCfg()
    cfg.env.num_envs = num_envs
    if mutate is not None:
        mutate(cfg)
    return cfg


def build_reference_legged_robot(state, seed=1, mutate_cfg=None, env_id_offset=0):
    """The UNMODIFIED ``LeggedRobot`` (legged_gym/envs/base/legged_robot.py) with the a1 config over the synthetic tensors
    of ``state`` (pointfoot_b200.synth.make_lr_state): built with ``object.__new__``, what BaseTask.__init__ / create_sim /
    _create_envs would have set from PhysX is set from the URDF constants, then the reference's own ``_parse_cfg``,
    ``_process_dof_props``, ``_get_env_origins``, ``_init_buffers`` and ``_prepare_reward_function`` run."""
    cls = _make_lr_class()
    from types import SimpleNamespace
    N = int(state["root_states"].shape[0])
    cfg = make_a1_cfg(N, mutate_cfg)
    env = object.__new__(cls)
    env.rng_seed, env.env_id_offset = seed, env_id_offset
    # ---- LeggedRobot.__init__ (LR:65-70) + BaseTask.__init__ (base_task.py:40-85) ----
    env.cfg = cfg
    env.sim_params = SimpleNamespace(dt=cfg.sim.dt, use_gpu_pipeline=False)
    env.height_samples = None
    env.debug_viz = False
    env.init_done = False
    env._parse_cfg(cfg)
    env.gym = FakeGym(state)
    env.sim = env.physics_engine = None
    env.sim_device = env.device = "cpu"
    env.headless, env.viewer, env.enable_viewer_sync = True, None, True
    env.num_envs, env.num_obs = N, cfg.env.num_observations
    env.num_privileged_obs, env.num_actions = cfg.env.num_privileged_obs, cfg.env.num_actions
    env.obs_buf = torch.zeros(N, env.num_obs)
    env.rew_buf = torch.zeros(N)
    env.reset_buf = torch.ones(N, dtype=torch.long)
    env.episode_length_buf = torch.zeros(N, dtype=torch.long)
    env.time_out_buf = torch.zeros(N, dtype=torch.bool)
    env.privileged_obs_buf = None
    env.extras = {}
    # ---- create_sim / _create_envs (LR:228-244, 614-699) ----
    env.up_axis_idx = 2
    env.num_dof = env.num_dofs = 12
    env.num_bodies = len(A1_BODY_NAMES)
    env.dof_names = list(A1_DOF_NAMES)
    props = np.zeros(12, dtype=[("lower", "f4"), ("upper", "f4"), ("velocity", "f4"), ("effort", "f4")])
    props["lower"], props["upper"], props["velocity"], props["effort"] = A1_DOF_LOWER, A1_DOF_UPPER, A1_DOF_VELOCITY, A1_DOF_EFFORT
    env._process_dof_props(props, 0)                         # reference code: soft limits, LR:279-305

    def idx(names):
        out = []
        for n in names:
            out.extend(i for i, s in enumerate(A1_BODY_NAMES) if n in s)
        return torch.tensor(out, dtype=torch.long)
    env.feet_indices = idx([cfg.asset.foot_name])
    env.penalised_contact_indices = idx(cfg.asset.penalize_contacts_on)
    env.termination_contact_indices = idx(cfg.asset.terminate_after_contacts_on)
    isl = cfg.init_state
    env.base_init_state = _tu.to_torch(isl.pos + isl.rot + isl.lin_vel + isl.ang_vel)
    if cfg.terrain.mesh_type in ("heightfield", "trimesh"):
        env.terrain = SimpleNamespace(cfg=cfg.terrain, env_length=cfg.terrain.terrain_length,
                                      env_origins=state["terrain_origins"].cpu().numpy())
        env.height_samples = state["height_samples"]
    env._get_env_origins()                                   # reference code, LR:701-726 (its torch.randint levels are overwritten)
    if env.custom_origins:
        env.terrain_levels = state["terrain_levels"].clone()
        env.terrain_types = state["terrain_types"].clone()
        env.env_origins[:] = env.terrain_origins[env.terrain_levels, env.terrain_types]
    env.envs, env.actor_handles = [None] * N, [None] * N
    env._init_buffers()
    env._prepare_reward_function()
    env.init_done = True
    # ---- internal state the synthetic generator randomises ----
    env.commands[:] = state["commands"]
    env.episode_length_buf[:] = state["episode_length_buf"]
    for k in ("last_actions", "last_dof_vel", "feet_air_time", "last_contacts"):
        getattr(env, k)[:] = state[k]
    keys = [k for k in env.episode_sums if k != "termination"] + (["termination"] if "termination" in env.episode_sums else [])
    for i, name in enumerate(keys):
        env.episode_sums[name][:] = state["episode_sums"][i]
    env.common_step_counter = int(state.get("common_step_counter", 0))
    return env


# --------------------------------------------------------------------------------------------
# Terrain generation: the UNMODIFIED legged_gym/utils/terrain.py::Terrain over the restated
# isaacgym.terrain_utils (oracle/terrain_ref.py -- third-party, absent, parity unpinned for that part).
def make_terrain_cfg(mutate=None):
    install_stubs()
    from legged_gym.envs.base.legged_robot_config import LeggedRobotCfg

    class terrain(LeggedRobotCfg.terrain):
        pass
    cfg = terrain()
    if mutate is not None:
        mutate(cfg)
    return cfg


def build_reference_terrain(mutate=None, seed=1, num_robots=64):
    """-> the reference's Terrain object (height_field_raw int16, env_origins float64[rows, cols, 3], ...).  Every random draw
    -- the primitives' through terrain_ref.draw_index, Terrain.randomized_terrain's own np.random.uniform / np.random.choice
    (terrain.py:78-79) -- is routed to the counter-based contract: purpose TERRAIN, 'env' = sub-terrain index
    row * num_cols + col, step 0."""
    install_stubs()
    from oracle import terrain_ref
    import legged_gym.envs.base.legged_robot_config                # (envs first: legged_gym.utils <-> envs import cycle)
    import legged_gym.utils.terrain as ref_terrain
    ref_terrain.terrain_utils = terrain_ref                  # the module global the reference's calls resolve through
    cfg = make_terrain_cfg(mutate)
    ctx = {"sub": 0, "made": 0, "layout": 0}

    def sub_of(count):
        if cfg.curriculum:                                   # terrain.py:85-92: for j: for i
            j, i = divmod(count, cfg.num_rows)
        else:                                                # terrain.py:73-82 / 94-107: k -> unravel_index
            i, j = divmod(count, cfg.num_cols)
        return i * cfg.num_cols + j

    class RefTerrain(ref_terrain.Terrain):
        def make_terrain(self, choice, difficulty):
            ctx["sub"] = sub_of(ctx["made"])
            ctx["made"] += 1
            return super().make_terrain(choice, difficulty)

    def hook(n, lane):
        return int(rng.rand_u32(seed, 0, rng.TERRAIN, ctx["sub"], lane)) % n

    def np_uniform(lo=0.0, hi=1.0, size=None):
        assert size is None
        sub = sub_of(ctx["layout"])
        return lo + (hi - lo) * float(rng.uniform(seed, 0, rng.TERRAIN, sub, terrain_ref.LANE_LAYOUT))

    def np_choice(a, size=None, **kw):
        assert size is None and not kw
        sub = sub_of(ctx["layout"])
        ctx["layout"] += 1
        return a[int(rng.rand_u32(seed, 0, rng.TERRAIN, sub, terrain_ref.LANE_LAYOUT + 1)) % len(a)]

    terrain_ref.set_draw_hook(hook)
    try:
        with _patched(np.random, "uniform", np_uniform), _patched(np.random, "choice", np_choice):
            return RefTerrain(cfg, num_robots)
    finally:
        terrain_ref.set_draw_hook(None)
