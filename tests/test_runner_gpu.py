"""The drop-in surface end to end on the GPU: task_registry.make_env / make_alg_runner / OnPolicyRunner.learn /
save / load / get_inference_policy (legged_gym/utils/task_registry.py:65-155, scripts/train.py:41-44, play.py:59-79),
and a closed-loop parity run of BASELINE.json configs[0] (pointfoot_flat, 64 envs, one PPO iteration) against the
oracle env + oracle rsl_rl."""
import os

import pytest
import torch

import common
import parity_rl
from pointfoot_b200 import config, helpers, synth

pytestmark = pytest.mark.gpu


def _args(task, n, iters=2):
    return helpers.get_args(["--task", task, "--num_envs", str(n), "--max_iterations", str(iters), "--headless"])


@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
def test_train_two_iterations_save_load(task, tmp_path, monkeypatch):
    monkeypatch.setenv("PF_RL_PRECISION", "bf16")
    from pointfoot_b200.task_registry import task_registry
    args = _args(task, 512)
    env, env_cfg = task_registry.make_env(name=task, args=args)
    runner, train_cfg = task_registry.make_alg_runner(env=env, name=task, args=args, log_root=str(tmp_path))
    assert env.num_envs == 512 and runner.alg.storage.observations.shape == (24, 512, 27)
    runner.learn(num_learning_iterations=2, init_at_random_ep_len=True)
    vl, sl = runner.last_losses
    assert torch.isfinite(torch.tensor([vl, sl])).all()
    sd = runner.alg.actor_critic.state_dict()
    assert {"std", "actor.0.weight", "actor.6.bias", "critic.0.weight", "critic.6.weight"} <= set(sd)
    run_dir = os.path.join(str(tmp_path), sorted(os.listdir(str(tmp_path)))[-1])
    path = helpers.get_load_path(str(tmp_path))
    assert path == os.path.join(run_dir, "model_2.pt")
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    assert set(ckpt) == {"model_state_dict", "optimizer_state_dict", "iter", "infos"} and ckpt["iter"] == 2
    # the checkpoint loads into the reference stack's module (the oracle restatement) unchanged ...
    from oracle import rsl_rl_ref
    nc = env.num_privileged_obs or env.num_obs
    oac = rsl_rl_ref.ActorCritic(env.num_obs, nc, env.num_actions, **config.class_to_dict(train_cfg)["policy"])
    oac.load_state_dict(ckpt["model_state_dict"])
    opt = torch.optim.Adam(oac.parameters(), lr=1e-3)
    opt.load_state_dict(ckpt["optimizer_state_dict"])
    # ... and back into ours; the inference policy then agrees with the torch module on the same weights
    before = runner.alg.actor_critic.flat.clone()
    runner.alg.actor_critic.flat.zero_()
    runner.load(path)
    assert torch.equal(runner.alg.actor_critic.flat, before)
    policy = runner.get_inference_policy(device=env.device)
    obs = env.get_observations()
    with torch.no_grad():
        want = oac.act_inference(obs.cpu())
    torch.testing.assert_close(policy(obs).cpu(), want, **parity_rl.BF16)
    runner.alg.close(); env.close()


def test_closed_loop_config_c1_against_oracle():
    """configs[0]: pointfoot_flat, 64 envs, one PPO iteration: env.step and PPO.act feed each other for 24 steps,
    then GAE and 5 x 4 minibatch updates, fp32 path.  Storage after the rollout and parameters after the update are
    compared with the oracle stack run on the same synthetic state."""
    from oracle import rsl_rl_ref
    from pointfoot_b200.env import PointFoot
    from pointfoot_b200 import rl
    N, T, task = 64, 24, "pointfoot_flat"
    state = synth.make_state(N, task, seed=1)
    cfg = common.make_cfg(task, N)
    tcfg = config.class_to_dict(config.PointFootFlatCfgPPO())
    env = PointFoot(cfg, None, None, "cuda:0", True, state=common.clone_state(state), seed=1)
    orc, _ = common.oracle_env(task, state)
    # the class starts from zeroed buffers like the reference's _init_buffers; give it the oracle's mid-episode state
    for k in ("episode_length_buf", "last_actions", "last_dof_vel", "feet_air_time", "last_feet_air_time",
              "current_max_feet_height", "last_max_feet_height", "contact_filt", "first_contact", "last_contacts"):
        getattr(env, k).copy_(state[k].view(getattr(env, k).shape))
    env._episode_sums.copy_(state["episode_sums"][:env._episode_sums.shape[0]])
    ppo, oppo = parity_rl.make_pair(N, 27, 27, tcfg["policy"]["actor_hidden_dims"], T, seed=1, alg=tcfg["algorithm"])
    try:
        obs, oobs = env.get_observations(), torch.zeros(N, 27)
        # both sides start from the state tensors as they are (no reset), zero first observation
        obs.zero_()
        for t in range(T):
            a = ppo.act(obs, obs)
            with torch.no_grad():
                oa = oppo.act(oobs, oobs)
            obs, _, rew, dones, infos = env.step(a)
            oobs, _, orew, odones = orc.step(oa.clone())
            ppo.process_env_step(rew, dones, infos)
            oppo.process_env_step(orew, odones, {"time_outs": orc.extras_time_outs})
            assert torch.equal(dones.cpu(), odones), f"reset flags diverged at step {t}"
            torch.testing.assert_close(obs.cpu(), oobs, rtol=1e-3, atol=1e-3, msg=lambda m, t=t: f"obs step {t}: {m}")
        parity_rl.check_storage(ppo, oppo, dict(rtol=2e-3, atol=2e-3), names=("actions", "values", "actions_log_prob", "rewards", "dones"))
        ppo.compute_returns(obs)
        with torch.no_grad():
            oppo.compute_returns(oobs)
        torch.testing.assert_close(ppo.storage.returns.cpu(), oppo.storage.returns, rtol=5e-3, atol=5e-3)
        p0 = parity_rl.flat_params(oppo.actor_critic).clone()
        ppo.update(); oppo.update()
        p1, q1 = parity_rl.flat_params(oppo.actor_critic), ppo.actor_critic.flat.cpu()
        assert (q1 - p1).norm() <= 0.1 * (p1 - p0).norm(), ((q1 - p1).norm(), (p1 - p0).norm())
    finally:
        ppo.close(); env.close()


def test_play_loop_and_jit_export(tmp_path):
    """scripts/play.py (reference play.py:42-114, headless): closed loop with the inference policy, TorchScript export of
    the actor (helpers.py:180-190); the exported CPU module agrees with the CUDA act_inference on the same weights."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("pf_play", os.path.join(common.ROOT, "scripts", "play.py"))
    play = importlib.util.module_from_spec(spec); spec.loader.exec_module(play)
    args = _args("pointfoot_rough", 32)
    file = play.play(args, steps=8, export_dir=str(tmp_path))
    assert file and os.path.isfile(file)
    m = torch.jit.load(file)
    from pointfoot_b200.task_registry import task_registry
    env, _ = task_registry.make_env(name="pointfoot_rough", args=args)
    runner, _ = task_registry.make_alg_runner(env=env, name="pointfoot_rough", args=args, log_root=None)
    ac = runner.alg.actor_critic
    x = torch.randn(7, 27)
    from pointfoot_b200 import helpers as h
    want = h.actor_module_from_state_dict(ac.state_dict())(x)
    got = ac.act_inference(x.to(env.device)).cpu()
    prec = parity_rl.BF16 if os.environ.get("PF_RL_PRECISION", "") == "bf16" else dict(rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(got, want, **prec)
    assert m(x).shape == (7, 6)
    runner.alg.close(); env.close()
