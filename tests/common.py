"""Shared helpers for the tests: build parameter blocks, oracle envs and buffer sets from one synthetic state."""
import copy
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from pointfoot_b200 import capi, config, params as pf_params, synth   # noqa: E402

CFGS = {"pointfoot_flat": config.PointFootFlatCfg, "pointfoot_rough": config.PointFootRoughCfg}
LR_GRID_X = [-0.8, -0.7, -0.6, -0.5, -0.4, -0.3, -0.2, -0.1, 0., 0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8]  # legged_robot_config.py:54
LR_GRID_Y = [-0.5, -0.4, -0.3, -0.2, -0.1, 0., 0.1, 0.2, 0.3, 0.4, 0.5]                                    # legged_robot_config.py:55


def make_cfg(task, num_envs, mutate=None):
    cfg = CFGS[task]()
    cfg.env.num_envs = num_envs
    if mutate:
        mutate(cfg)
    return cfg


def grid187(cfg):
    """the 17x11 = 187-point LeggedRobotCfg scan grid"""
    cfg.terrain.measured_points_x = list(LR_GRID_X)
    cfg.terrain.measured_points_y = list(LR_GRID_Y)
    cfg.env.num_privileged_obs = 27 + 187


def all_reward_terms(cfg):
    """Every optional branch of compute_reward (PF:301-318): the five terms both pointfoot cfgs leave at -0.0
    (PF:1153 dof_vel, :1170 termination, :1174 dof_pos_limits, :1180 dof_vel_limits, :1226 stumble) get non-zero scales
    and ``only_positive_rewards`` is switched on (the clip of PF:312-313 and the post-clip termination add of
    PF:315-318).  The cfg key is ``stumble``: ``feet_stumble`` has no ``_reward_feet_stumble``."""
    sc = cfg.rewards.scales
    sc.dof_vel = -2.e-4
    sc.termination = -4.0
    sc.dof_pos_limits = -3.0
    sc.dof_vel_limits = -0.7
    sc.stumble = -1.5
    sc.feet_stumble = 0.0
    # tracking terms dominate less than half of the envs, so the clip hits some envs and not others
    sc.tracking_lin_vel = 12.0
    cfg.rewards.only_positive_rewards = True


def control_v(cfg):
    cfg.control.control_type = "V"          # PF:540-542


def control_t(cfg):
    cfg.control.control_type = "T"          # PF:543-544


def push_every_step_2(cfg):
    """push at common_step_counter 2 (and 4): push_interval = ceil(push_interval_s / dt) = 2 (PF:1009, 492-493)"""
    cfg.domain_rand.push_robots = True
    cfg.domain_rand.push_interval_s = 0.04


def clone_state(state):
    return {k: (v.clone() if torch.is_tensor(v) else copy.deepcopy(v)) for k, v in state.items()}


def make_params(task, state, mutate=None, seed=1, env_id_offset=0, index_div_mode=capi.DIV_TRUE,
                resample_commands=False):
    N = state["root_states"].shape[0]
    cfg = make_cfg(task, N, mutate)
    hf = tuple(state["height_samples"].shape) if "height_samples" in state else (0, 0)
    return (cfg,) + pf_params.build_params(cfg, N, seed=seed, env_id_offset=env_id_offset, hf_shape=hf,
                                           index_div_mode=index_div_mode, resample_commands=resample_commands)


def alloc_buffers(P, state, device="cpu", with_measured_heights=True):
    """Tensors for every pf_env_buffers field, initialised from the synthetic state.  Returns (dict, PfEnvBuffers)."""
    N, D, R = P.num_envs, P.num_dof, P.num_reward_terms + P.has_termination_term
    dev = torch.device(device)
    z = lambda *s, dt=torch.float: torch.zeros(*s, dtype=dt, device=dev)
    t = {}
    for k in ("root_states", "dof_state", "contact_forces", "rigid_body_states", "commands", "episode_length_buf",
              "last_actions", "last_dof_vel", "feet_air_time", "last_feet_air_time", "current_max_feet_height",
              "last_max_feet_height", "contact_filt", "first_contact", "last_contacts"):
        t[k] = state[k].clone().to(dev).contiguous()
    t["episode_sums"] = state["episode_sums"][:R].clone().to(dev).contiguous()
    t["actions"], t["torques"] = z(N, D), z(N, D)
    t["base_lin_vel"], t["base_ang_vel"], t["projected_gravity"] = z(N, 3), z(N, 3), z(N, 3)
    t["last_root_vel"], t["feet_height"] = z(N, 6), z(N, 2)
    t["rew_buf"] = z(N)
    t["reset_buf"] = torch.ones(N, dtype=torch.bool, device=dev)
    t["time_out_buf"], t["terminate"] = z(N, dt=torch.bool), z(N, dt=torch.bool)
    t["obs_buf"] = z(N, P.num_obs)
    t["priv_obs_buf"] = z(N, P.num_priv_obs) if P.num_priv_obs else None
    t["measured_heights"] = z(N, P.num_height_points) if (P.num_height_points and with_measured_heights) else None
    if P.terrain_kind == capi.TERRAIN_HEIGHTFIELD:
        t["height_samples"] = state["height_samples"].to(dev).contiguous()
        t["terrain_origins"] = state["terrain_origins"].to(dev).contiguous()
        t["terrain_levels"] = state["terrain_levels"].clone().to(dev)
        t["terrain_types"] = state["terrain_types"].clone().to(dev)
        t["env_origins"] = t["terrain_origins"][t["terrain_levels"], t["terrain_types"]].clone().contiguous()
    else:
        from oracle.pointfoot_ref import grid_env_origins
        t["height_samples"] = t["terrain_origins"] = t["terrain_levels"] = t["terrain_types"] = None
        t["env_origins"] = (state["env_origins"].clone() if "env_origins" in state else grid_env_origins(N, 3.0)).to(dev)
    t["reset_ids"] = torch.zeros(N, dtype=torch.int32, device=dev)
    t["reset_count"] = torch.zeros(1, dtype=torch.int32, device=dev)
    t["episode_stats"] = z(R + 2)
    t["extras_time_outs"] = z(N, dt=torch.bool)
    t["step_counter"] = None
    b = capi.PfEnvBuffers()
    for name in capi.ENV_BUFFER_FIELDS:
        setattr(b, name, None if t[name] is None else t[name].data_ptr())
    return t, b


_emu = None


def host_emu():
    """ctypes handle of the host emulation of the kernels (tests/host_emu)."""
    global _emu
    if _emu is None:
        subprocess.check_call([os.path.join(ROOT, "tests/host_emu/build.sh")])
        _emu = C.CDLL(os.path.join(ROOT, "tests/host_emu/_build/libpf_host_emu.so"))
    return _emu


def oracle_env(task, state, mutate=None, **kw):
    from oracle.pointfoot_ref import RefEnv
    cfg, P, noise, pts, names = make_params(task, state, mutate, **kw)
    orc = RefEnv(P, noise, pts, clone_state(state))
    orc.push_interval, orc.sim_dt_py, orc.max_push_vel_xy_py = push_schedule(cfg)
    return orc, (cfg, P, noise, pts, names)


def push_schedule(cfg):
    """(push_interval in policy steps or 0, sim dt, max_push_vel_xy) as the reference holds them (PF:492-493, 592-596, 1009)"""
    dt = cfg.control.decimation * cfg.sim.dt
    interval = int(np.ceil(cfg.domain_rand.push_interval_s / dt)) if cfg.domain_rand.push_robots else 0
    return interval, cfg.sim.dt, cfg.domain_rand.max_push_vel_xy


class RedZone:
    """Guard bands around device buffers: every tensor handed to a kernel is re-homed in the middle of a larger allocation
    filled with 0xA5, and the bands are checked afterwards -- catches out-of-bounds WRITES of either sign (the stand-in
    for compute-sanitizer memcheck, which is closed on this GPU pool: profiles/r2_sanitizer_unavailable.txt)."""
    PAD = 1024

    def __init__(self):
        self.guards = []

    def wrap(self, t, name=""):
        n = t.numel() * t.element_size()
        raw = torch.full((n + 2 * self.PAD,), 0xA5, dtype=torch.uint8, device=t.device)
        mid = raw[self.PAD:self.PAD + n]
        mid.copy_(t.contiguous().view(-1).view(torch.uint8))
        self.guards.append((name, raw, n))
        return mid.view(t.dtype).view(t.shape)

    def check(self):
        for name, raw, n in self.guards:
            lo, hi = raw[:self.PAD], raw[self.PAD + n:]
            assert bool((lo == 0xA5).all()), f"write BELOW buffer {name}: {int((lo != 0xA5).sum())} bytes"
            assert bool((hi == 0xA5).all()), f"write ABOVE buffer {name}: {int((hi != 0xA5).sum())} bytes"
        return len(self.guards)
