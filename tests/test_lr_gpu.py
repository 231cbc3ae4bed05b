"""LeggedRobot / a1 on the GPU, through the C-ABI (pf_lr_pd_torques / pf_lr_post_physics_step /
pf_lr_reset_all): against the golden outputs of the UNMODIFIED reference class, and -- at sizes the golden files do not
cover -- against the host emulation of the same per-env code, which the CPU suite pins to those golden outputs."""
import pytest
import torch

import parity_lr

pytestmark = pytest.mark.gpu

STATE_BUFFERS = ("obs_buf", "rew_buf", "reset_buf", "time_out_buf", "episode_length_buf", "root_states", "dof_state", "commands",
                 "torques", "actions", "last_actions", "last_dof_vel", "last_root_vel", "feet_air_time", "last_contacts",
                 "base_lin_vel", "base_ang_vel", "projected_gravity", "measured_heights", "terrain_levels", "env_origins",
                 "_episode_sums", "_time_outs")
# values that pass through expf / atan2f / a 12-term or 187-term sum: CUDA's libm and nvcc's loop order against glibc / g++
CLOSE = {"obs_buf", "rew_buf", "_episode_sums", "commands"}


@pytest.mark.parametrize("fname", list(parity_lr.LR_GOLDEN))
def test_cuda_path_matches_reference_golden(fname):
    parity_lr.golden_impl(fname, "cuda:0")


def _pair(n, mutate, seed, start):
    from pointfoot_b200 import config, synth
    from pointfoot_b200.legged_robot import LeggedRobot
    state = synth.make_lr_state(n, seed=seed, adversarial=True)
    envs = []
    for cls, dev in ((LeggedRobot, "cuda:0"), (parity_lr.emu_class(), "cpu")):
        cfg = config.A1RoughCfg()
        cfg.env.num_envs = n
        if mutate is not None:
            mutate(cfg)
        env = cls(cfg, None, None, dev, True, state={k: (v.clone() if torch.is_tensor(v) else v) for k, v in state.items()}, seed=seed)
        env.common_step_counter = start
        envs.append(env)
    return envs


def _compare(gpu, emu, label):
    for k in STATE_BUFFERS:
        a, b = getattr(gpu, k), getattr(emu, k)
        if not torch.is_tensor(b):
            continue
        a = a.cpu()
        if k in CLOSE:
            torch.testing.assert_close(a, b, rtol=parity_lr.RTOL, atol=parity_lr.ATOL, msg=lambda m: f"[{label}] {k}: {m}")
        else:
            assert torch.equal(a, b), f"[{label}] {k} not bit-exact: {int((a != b).sum())} of {a.numel()}"
    n = int(gpu._reset_count.item())
    assert n == int(emu._reset_count.item())
    assert torch.equal(gpu._reset_ids[:n].cpu(), emu._reset_ids[:n]), f"[{label}] reset ids"
    if n:
        torch.testing.assert_close(gpu._episode_stats.cpu(), emu._episode_stats, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n,mutate,start", [(4096, None, 748), (33, None, 0), (1, None, 0), (5000, parity_lr.lr_flat_all_terms, 0),
                                            (2049, parity_lr.lr_rough_no_noise_base_height, 0)])
def test_cuda_path_matches_host_emulation_at_other_sizes(n, mutate, start):
    gpu, emu = _pair(n, mutate, seed=3, start=start)
    for i in range(3):
        a = parity_lr.actions_for(n, 3, i)
        gpu.step(a.cuda())
        emu.step(a)
        torch.cuda.synchronize()
        _compare(gpu, emu, f"N={n} step {i}")


def test_reset_matches_host_emulation():
    """base_task.py:105-109 reset(): reset_idx over every env, then a zero-action step"""
    gpu, emu = _pair(777, None, seed=5, start=10)
    og, _ = gpu.reset()
    oe, _ = emu.reset()
    assert int(gpu.episode_length_buf.max()) <= 1
    torch.testing.assert_close(og.cpu(), oe, rtol=parity_lr.RTOL, atol=parity_lr.ATOL)
    _compare(gpu, emu, "reset")


def test_rejects_a_mismatched_observation_width():
    from pointfoot_b200 import config
    from pointfoot_b200.legged_robot import LeggedRobot
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = 16
    cfg.env.num_observations = 48          # but the height scan is on: LR:224 would concatenate 235 columns
    env = LeggedRobot(cfg, None, None, "cuda:0", True)
    with pytest.raises(RuntimeError, match="observation width"):
        env.step(torch.zeros(16, 12, device="cuda"))
