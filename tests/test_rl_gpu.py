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
