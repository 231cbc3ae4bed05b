"""Oracle pinning (CPU): the plain-torch restatement and the host emulation of the kernels' per-env logic both
reproduce the outputs of the UNMODIFIED reference point_foot.py stored in tests/golden (made by make_golden.py)."""
import pytest

import parity


@pytest.mark.parametrize("fname", sorted(parity.GOLDEN))
def test_oracle_matches_reference_golden(fname):
    parity.golden_oracle(fname)


@pytest.mark.parametrize("fname", sorted(parity.GOLDEN))
def test_kernel_core_host_emulation_matches_reference_golden(fname):
    parity.golden_impl(fname, device="cpu", backend="emu")


@pytest.mark.parametrize("task,n", [("pointfoot_flat", 1), ("pointfoot_flat", 33), ("pointfoot_rough", 257)])
def test_host_emulation_matches_oracle(task, n):
    parity.run(task, n, steps=3, device="cpu", backend="emu")


def test_live_reference_matches_golden():
    """where /root/reference exists, re-run the unmodified reference and compare with the committed fixture"""
    from oracle import ref_harness
    if not ref_harness.reference_available():
        pytest.skip("reference tree not present (GPU box)")
    import common
    for fname in sorted(parity.GOLDEN):
        task, mutate, state, _ = parity.load_golden(fname)
        env = ref_harness.build_reference_env(task, common.clone_state(state), mutate_cfg=mutate)   # (state carries common_step_counter)
        last = {}

        def stepper(i, a):
            last["out"] = env.step(a.clone())

        def fetch(name):
            if name == "reward_names":
                return list(env.episode_sums.keys())
            obs, priv, rew, reset, extras = last["out"]
            special = {"extras_time_outs": extras["time_outs"], "obs_buf": obs, "priv_obs_buf": priv, "rew_buf": rew, "reset_buf": reset,
                       "reset_ids": reset.nonzero().flatten()}
            if name in special:
                return special[name]
            if name == "push_forces":
                return env.rigid_body_external_forces
            if name == "episode_sums":
                import torch
                return torch.stack([env.episode_sums[n] for n in env.episode_sums.keys()])
            if name == "episode_stats":
                import torch
                ep = extras["episode"]
                return torch.stack([ep["rew_" + n] for n in env.episode_sums.keys()])
            return getattr(env, name, None)
        parity.check_against_golden(fname, stepper, fetch)


@pytest.mark.parametrize("var", ["PF_EMU_GENERIC", "PF_EMU_TABLE"])
@pytest.mark.parametrize("task,n", [("pointfoot_flat", 33), ("pointfoot_rough", 65)])
def test_other_statements_of_the_step_match_oracle(task, n, var, monkeypatch):
    """the step is compiled three ways (pf_env_core.cuh, pf_canon_pos): the pointfoot term set (default), the
    table-driven sum for any subset in dir() order, the order-agnostic loop; the last two agree with the oracle too"""
    monkeypatch.setenv(var, "1")
    parity.run(task, n, steps=3, device="cpu", backend="emu")


@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
def test_host_emulation_matches_oracle_every_optional_branch(task):
    """all 24 reward terms + only_positive_rewards + termination add (PF:301-318, 1132-1253), a push step (PF:590-611),
    control types V / T (PF:540-545), at a ragged env count"""
    import common
    parity.run(task, 97, steps=3, device="cpu", backend="emu", mutate=common.all_reward_terms, num_reward_terms=24)
    parity.run(task, 97, steps=3, device="cpu", backend="emu", start_step=348)
    parity.run(task, 97, steps=2, device="cpu", backend="emu", mutate=common.control_v)
    parity.run(task, 97, steps=2, device="cpu", backend="emu", mutate=common.control_t)
