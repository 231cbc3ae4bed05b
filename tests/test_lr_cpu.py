"""LeggedRobot / a1, CPU suite: the per-env code of the pf_lr_* kernels (pf_lr_core.cuh, driven by
tests/host_emu) against the golden outputs of the UNMODIFIED reference class; the reference re-run live where it exists;
the host-side constants against the reference's own."""
import numpy as np
import pytest
import torch

import parity_lr
from oracle import ref_harness

needs_reference = pytest.mark.skipif(not ref_harness.reference_available(), reason="/root/reference not present")


@pytest.mark.parametrize("fname", list(parity_lr.LR_GOLDEN))
def test_kernel_core_host_emulation_matches_reference_golden(fname):
    parity_lr.golden_impl(fname, "cpu")


@needs_reference
@pytest.mark.parametrize("fname", list(parity_lr.LR_GOLDEN))
def test_live_reference_reproduces_golden(fname):
    parity_lr.golden_live_reference(fname)


@needs_reference
def test_a1_config_equals_the_references():
    from pointfoot_b200 import config
    ref = ref_harness.make_a1_cfg(64)
    mine = config.A1RoughCfg()
    for sect in ("env", "commands", "init_state", "control", "domain_rand", "rewards", "normalization", "noise"):
        a, b = config.class_to_dict(getattr(mine, sect)), config.class_to_dict(getattr(ref, sect))
        a.pop("num_envs", None); b.pop("num_envs", None)
        assert a == b, sect
    for k in ("mesh_type", "horizontal_scale", "vertical_scale", "border_size", "curriculum", "measure_heights",
              "measured_points_x", "measured_points_y", "terrain_length", "num_rows", "num_cols", "max_init_terrain_level"):
        assert getattr(mine.terrain, k) == getattr(ref.terrain, k), k
    for k in ("foot_name", "penalize_contacts_on", "terminate_after_contacts_on"):
        assert getattr(mine.asset, k) == getattr(ref.asset, k), k
    from legged_gym.envs.a1.a1_config import A1RoughCfgPPO
    for sect in ("policy", "algorithm"):
        assert config.class_to_dict(getattr(config.A1RoughCfgPPO(), sect)) == config.class_to_dict(getattr(A1RoughCfgPPO(), sect))
    r = config.class_to_dict(A1RoughCfgPPO().runner)
    m = config.class_to_dict(config.A1RoughCfgPPO().runner)
    assert {k: m[k] for k in r} == r


@needs_reference
def test_lr_params_equal_what_the_reference_prepares():
    """build_lr_params against the attributes the reference's own _parse_cfg / _process_dof_props / _init_buffers /
    _prepare_reward_function leave"""
    from pointfoot_b200 import config, legged_robot, synth
    state = synth.make_lr_state(8, seed=2)
    ref = ref_harness.build_reference_legged_robot({k: (v.clone() if torch.is_tensor(v) else v) for k, v in state.items()})
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = 8
    P, noise, pts, names = legged_robot.build_lr_params(cfg, 8, hf_shape=tuple(state["height_samples"].shape))
    arr = lambda a, n: np.array([a[i] for i in range(n)], np.float32)
    assert np.array_equal(noise.numpy(), ref.noise_scale_vec.numpy())
    assert np.array_equal(pts.numpy(), ref.height_points[0, :, :2].numpy())
    assert np.array_equal(arr(P.p_gains, 12), ref.p_gains.numpy()) and np.array_equal(arr(P.d_gains, 12), ref.d_gains.numpy())
    assert np.array_equal(arr(P.default_dof_pos, 12), ref.default_dof_pos[0].numpy())
    assert np.array_equal(arr(P.torque_limits, 12), ref.torque_limits.numpy())
    assert np.array_equal(np.array([[P.dof_pos_limits[i][0], P.dof_pos_limits[i][1]] for i in range(12)], np.float32),
                          ref.dof_pos_limits.numpy())
    assert [P.feet_indices[i] for i in range(P.num_feet)] == ref.feet_indices.tolist()
    assert [P.penalised_indices[i] for i in range(P.num_penalised)] == ref.penalised_contact_indices.tolist()
    assert [P.termination_indices[i] for i in range(P.num_termination)] == ref.termination_contact_indices.tolist()
    assert names[:P.num_reward_terms] == [n for n in ref.reward_names]
    assert np.array_equal(arr(P.reward_scales, P.num_reward_terms), np.array([ref.reward_scales[n] for n in ref.reward_names], np.float32))
    assert P.max_episode_length == float(ref.max_episode_length) and P.resample_period == int(ref.cfg.commands.resampling_time / ref.dt)
    assert legged_robot.LR_REWARD_NAMES and all(hasattr(ref, "_reward_" + n) for n in legged_robot.LR_REWARD_NAMES)


def test_unknown_reward_name_fails_like_the_reference():
    from pointfoot_b200 import config, legged_robot
    cfg = config.A1RoughCfg()
    cfg.rewards.scales.feet_stumble = -1.0            # the base cfg's name for a method that is called _reward_stumble
    with pytest.raises(AttributeError, match="_reward_feet_stumble"):
        legged_robot.build_lr_params(cfg, 4)


def test_legged_robot_has_no_cpu_fallback():
    from pointfoot_b200 import config
    from pointfoot_b200.legged_robot import LeggedRobot
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = 4
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        LeggedRobot(cfg, None, None, "cpu", True)
