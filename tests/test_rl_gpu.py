"""GPU parity of the rsl_rl side (ActorCritic act/evaluate, storage writes, GAE, minibatch shuffle, PPO.update)
through the C-ABI, against oracle/rsl_rl_ref.py.  Tolerances are stated in parity_rl.py."""
import pytest

import parity_rl

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("hidden,nc", [((128, 64, 32), 27), ((512, 256, 128), 148)])
def test_act_storage_and_returns_match_oracle(hidden, nc):
    parity_rl.run_act_and_gae(num_envs=300, num_critic_obs=nc, hidden=hidden, T=6)


@pytest.mark.parametrize("hidden,nc,n", [((128, 64, 32), 27, 300), ((512, 256, 128), 148, 300), ((512, 256, 128), 148, 4096),
                                         ((256, 128), 48, 129), ((64, 64, 64, 64), 235, 1),
                                         # more items than SMs: every persistent CTA walks several (net, tile) items
                                         ((512, 256, 128), 148, 40000), ((128, 64, 32), 27, 50001)])
def test_tcgen05_mlp_act_matches_oracle_bf16(hidden, nc, n):
    """PF_RL_BF16_TC: fused tcgen05 MLP (bf16 operands, fp32 accumulate in TMEM); tolerance rtol = atol = 2e-2"""
    parity_rl.run_act_and_gae(num_envs=n, num_critic_obs=nc, hidden=hidden, T=3, precision="bf16")


def test_gae_bit_exact():
    parity_rl.run_gae_exact(num_envs=1000, T=24)
    parity_rl.run_gae_exact(num_envs=1, T=3, seed=5)


@pytest.mark.parametrize("n,t,mb", [(4096, 24, 4), (64, 24, 4), (33, 7, 2), (1, 1, 1)])
def test_minibatch_permutation_matches_contract(n, t, mb):
    parity_rl.run_permutation(n, t, iteration=3, num_mini_batches=mb)


def test_ppo_update_flat_net():
    parity_rl.run_update(num_envs=256, num_critic_obs=27, hidden=(128, 64, 32), T=24)


def test_ppo_update_rough_net():
    parity_rl.run_update(num_envs=256, num_critic_obs=148, hidden=(512, 256, 128), T=24)


@pytest.mark.parametrize("hidden,nc,n", [((512, 256, 128), 148, 256), ((128, 64, 32), 27, 700)])
def test_ppo_update_tensor_core_gemms_bf16(hidden, nc, n):
    """PF_RL_BF16_TC: forward / dgrad / wgrad of PPO.update on the tcgen05 GEMM (bf16 operands, fp32 accumulate);
    first-minibatch gradient within 10 x rtol 2e-2 of the fp32 oracle (relative to the largest gradient entry),
    parameters after the whole update within 25 % of the update's own L2 movement."""
    parity_rl.run_update(num_envs=n, num_critic_obs=nc, hidden=hidden, T=24, precision="bf16")


def test_rollout_buffer_overflow_raises():
    ppo, oppo = parity_rl.make_pair(8, 27, 27, [32, 16, 8], T=2)
    try:
        parity_rl.rollout(ppo, oppo, 2)
        import torch
        with pytest.raises(AssertionError, match="Rollout buffer overflow"):
            ppo.act(torch.zeros(8, 27, device="cuda"), torch.zeros(8, 27, device="cuda"))
    finally:
        ppo.close()


@pytest.mark.parametrize("hidden,nc,n", [((128, 64, 32), 27, 300), ((512, 256, 128), 148, 4096), ((128, 64, 32), 27, 50001)])
def test_rehearsed_act_is_bit_identical(hidden, nc, n):
    """pf_rl_tc_rehearse: the dry item the fused kernel runs before its dependency wait changes no output bit
    (actions, values, log-probs, storage rows), with one item per CTA and with several."""
    import torch
    from pointfoot_b200 import rl
    outs = []
    for on in (0, 1):
        torch.manual_seed(11)
        ac = rl.ActorCritic(27, nc, 6, hidden, hidden)
        ppo = rl.PPO(ac, precision="bf16", num_mini_batches=4, num_learning_epochs=1)
        ppo.init_storage(n, 3, [27], [nc], [6])
        assert ppo._lib.pf_rl_tc_rehearse(ppo._ctx, on) == 0
        g = torch.Generator(device="cuda").manual_seed(5)
        res = []
        for t in range(3):
            obs = torch.randn(n, 27, device="cuda", generator=g)
            cobs = torch.randn(n, nc, device="cuda", generator=g)
            a = ppo.act(obs, cobs).clone()
            ppo.storage.step += 1 if ppo.storage.step == t else 0
            res.append(a)
        torch.cuda.synchronize()
        st = ppo.storage
        outs.append([torch.stack(res)] + [getattr(st, k).clone() for k in ("observations", "privileged_observations", "actions", "values",
                                                                           "actions_log_prob", "mu", "sigma")])
        ppo.close()
    for x, y in zip(*outs):
        assert torch.equal(x, y)
