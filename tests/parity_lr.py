"""Parity driver of the LeggedRobot (a1) step path, SURVEY.md section 8f row N2.  TEST INFRASTRUCTURE.

The checker is the UNMODIFIED reference class (legged_gym/envs/base/legged_robot.py behind oracle/ref_harness.py): its
outputs on seeded synthetic states are committed as tests/golden/a1_*.npz (tests/golden/make_golden.py::generate_lr), and
where /root/reference exists the CPU suite also re-runs it live.  The implementation under test is
pointfoot_b200.legged_robot.LeggedRobot, i.e. the pf_lr_* entry points of the C-ABI.

Bars: integer / boolean / index outputs bit-exact; fp32 state that involves no libm call and no long sum bit-exact; the
rest (observations: atan2f in the heading command; rewards: expf, 12-term sums; episode sums) within RTOL / ATOL.
"""
import hashlib
import os

import numpy as np
import torch

from pointfoot_b200 import synth

RTOL, ATOL = 1e-5, 2e-6
GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STEPS = 3


# ---- cfg mutations (applied to the reference's A1RoughCfg when generating, to pointfoot_b200.config.A1RoughCfg when testing)
def lr_flat_all_terms(cfg):
    """plane, no height scan (48 observations), yaw-rate commands, velocity control, every _reward_* the class defines,
    no clip of the total, a push every 2nd step"""
    cfg.terrain.mesh_type = "plane"
    cfg.terrain.measure_heights = False
    cfg.terrain.curriculum = False
    cfg.env.num_observations = 48
    cfg.commands.heading_command = False
    cfg.control.control_type = "V"
    cfg.rewards.only_positive_rewards = False
    sc = cfg.rewards.scales
    sc.termination, sc.orientation, sc.base_height, sc.dof_vel, sc.stand_still = -200.0, -0.5, -1.0, -1e-4, -0.1
    sc.stumble, sc.feet_contact_forces, sc.dof_vel_limits, sc.torque_limits = -0.3, -0.01, -0.2, -0.05
    cfg.rewards.soft_dof_vel_limit, cfg.rewards.soft_torque_limit, cfg.rewards.max_contact_force = 0.1, 0.05, 30.0
    cfg.domain_rand.push_interval_s = 0.04


def lr_rough_no_noise_base_height(cfg):
    """rough terrain with the scan, noise off, base_height on (the 187-term mean), torque control, short episodes"""
    cfg.noise.add_noise = False
    cfg.control.control_type = "T"
    cfg.rewards.scales.base_height = -2.0
    cfg.rewards.scales.termination = -5.0
    cfg.env.episode_length_s = 10
    cfg.commands.resampling_time = 5.0


# file -> (mutation, common_step_counter before the first step).  748: the 2nd step is a push step (750 % 750 == 0)
LR_GOLDEN = {
    "a1_rough_n64.npz": (None, 748),
    "a1_flat_allterms_n64.npz": (lr_flat_all_terms, 0),
    "a1_rough_nonoise_n64.npz": (lr_rough_no_noise_base_height, 3),
}

EXACT = ("reset", "time_out_buf", "episode_length_buf", "last_contacts", "terrain_levels", "time_outs")
EXACT_F32 = ("torques", "base_lin_vel", "base_ang_vel", "projected_gravity", "last_actions", "last_dof_vel", "last_root_vel",
             "feet_air_time", "root_states", "dof_state", "env_origins", "measured_heights")
CLOSE = ("obs", "rew", "episode_sums", "commands")
ATTR = {"obs": "obs_buf", "rew": "rew_buf", "reset": "reset_buf", "time_outs": "_time_outs", "episode_sums": "_episode_sums"}


def actions_for(n, seed, i):
    g = torch.Generator().manual_seed(seed * 7919 + 100 + i)
    return torch.randn(n, 12, generator=g) * (30.0 if i == 1 else 1.0)      # step 1 reaches clip_actions = 100


def load_golden(fname):
    mutate, start_step = LR_GOLDEN[fname]
    z = np.load(os.path.join(GOLDEN_DIR, fname))
    state = {k[3:]: torch.from_numpy(z[k].copy()) for k in z.files if k.startswith("in_")}
    state["common_step_counter"] = start_step
    assert int(z["start_step"]) == start_step
    hf, origins = synth.make_heightfield(1)
    assert hashlib.sha256(hf.numpy().tobytes()).hexdigest() == bytes(z["hf_sha256"]).decode(), \
        "synthetic heightfield drifted from the one the golden run used"
    state["height_samples"] = hf
    return mutate, state, z


def check_steps(fname, z, step, fetch, names, steps=STEPS):
    """``step(i, actions)`` advances the implementation, ``fetch(attr)`` returns a tensor.  The reference keeps episode sums
    as a dict in dir() order with 'termination' at its alphabetical place; the C-ABI rows are matched BY NAME."""
    n = z["in_root_states"].shape[0]
    ref_names = [str(x) for x in z["reward_names"]]
    assert sorted(ref_names) == sorted(names), "enabled reward terms differ from the reference's"
    rows = [ref_names.index(x) for x in names]
    for i in range(steps):
        step(i, actions_for(n, 1, i))
        p = f"s{i}_"
        for name in EXACT + EXACT_F32 + CLOSE:
            if p + name not in z.files:
                continue
            got = fetch(ATTR.get(name, name))
            if got is None or not torch.is_tensor(got):
                continue
            want = torch.from_numpy(z[p + name])
            if name == "episode_sums":
                want = want[rows]
            got = got.detach().cpu().reshape(want.shape).to(want.dtype)
            if name in CLOSE:
                torch.testing.assert_close(got, want, rtol=RTOL, atol=ATOL, msg=lambda m, k=name: f"[{fname} step {i}] {k}: {m}")
            else:
                bad = (got != want) & ~(torch.isnan(got) & torch.isnan(want)) if got.is_floating_point() else got != want
                assert not bad.any(), f"[{fname} step {i}] {name} not bit-exact vs the reference: {int(bad.sum())} of {bad.numel()}"
        cnt = int(fetch("_reset_count").item())
        ids = fetch("_reset_ids")[:cnt].cpu().to(torch.int32)
        assert torch.equal(ids, torch.from_numpy(z[p + "reset_ids"])), f"[{fname} step {i}] reset ids"
        want = z[p + "extras_episode"]
        if want.size and cnt:
            stats = fetch("_episode_stats").cpu()
            torch.testing.assert_close(stats[:want.size], torch.from_numpy(want)[rows], rtol=1e-5, atol=1e-6)
            if p + "extras_terrain_level" in z.files:
                torch.testing.assert_close(stats[want.size], torch.tensor(float(z[p + "extras_terrain_level"])), rtol=1e-6, atol=0)


def emu_class():
    """LeggedRobot with its three C-ABI calls redirected to the host emulation of the same per-env code (tests/host_emu):
    lets the CPU suite check pf_lr_core.cuh against the reference's golden outputs without a GPU"""
    import ctypes as C
    import common
    from pointfoot_b200.legged_robot import LeggedRobot

    class EmuLeggedRobot(LeggedRobot):
        def _open_backend(self):
            self._lib = common.host_emu()

        def _call_pd_torques(self, actions):
            assert self._lib.pf_emu_lr_pd_torques(*self._args(), None if actions is None else C.c_void_p(actions.data_ptr())) == 0

        def _call_post_physics_step(self, push):
            assert self._lib.pf_emu_lr_post_physics_step(*self._args(), C.c_longlong(self.common_step_counter), C.c_int(push)) == 0

        def _call_reset_all(self):
            assert self._lib.pf_emu_lr_reset_all(*self._args(), C.c_longlong(self.common_step_counter)) == 0
    return EmuLeggedRobot


def make_impl(fname, device="cuda:0"):
    """pointfoot_b200.legged_robot.LeggedRobot over the golden file's input state (device 'cpu': the host emulation)"""
    from pointfoot_b200 import config
    from pointfoot_b200.legged_robot import LeggedRobot
    if device == "cpu":
        LeggedRobot = emu_class()
    mutate, state, z = load_golden(fname)
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = state["root_states"].shape[0]
    if mutate is not None:
        mutate(cfg)
    env = LeggedRobot(cfg, None, None, device, True, state={k: (v.clone() if torch.is_tensor(v) else v) for k, v in state.items()})
    env.common_step_counter = int(state["common_step_counter"])
    return env, z


def golden_impl(fname, device="cuda:0"):
    env, z = make_impl(fname, device)
    try:
        check_steps(fname, z, lambda i, a: env.step(a.to(device)), lambda k: getattr(env, k, None), list(env.reward_names))
    finally:
        env.close()


def golden_live_reference(fname):
    """the unmodified reference, re-run here, reproduces the committed file (only where /root/reference exists)"""
    from oracle import ref_harness
    mutate, state, z = load_golden(fname)
    env = ref_harness.build_reference_legged_robot({k: (v.clone() if torch.is_tensor(v) else v) for k, v in state.items()},
                                                   seed=1, mutate_cfg=mutate)
    n = state["root_states"].shape[0]
    for i in range(STEPS):
        obs, _, rew, reset, _ = env.step(actions_for(n, 1, i))
        assert np.array_equal(obs.numpy(), z[f"s{i}_obs"]) and np.array_equal(rew.numpy(), z[f"s{i}_rew"])
        assert np.array_equal(reset.numpy(), z[f"s{i}_reset"])
        assert np.array_equal(env.root_states.numpy(), z[f"s{i}_root_states"])
