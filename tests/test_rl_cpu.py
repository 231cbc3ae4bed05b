"""CPU checks of the rsl_rl oracle restatement (oracle/rsl_rl_ref.py) and of the host-side contracts."""
import numpy as np
import pytest
import torch

from oracle import rsl_rl_ref


@pytest.mark.parametrize("n", [1, 2, 5, 64, 1000, 98304])
def test_feistel_permutation_is_a_bijection(n):
    p = rsl_rl_ref.feistel_permutation(n, seed=1, iteration=2)
    assert p.shape == (n,) and np.array_equal(np.sort(p), np.arange(n))
    if n > 64:
        assert not np.array_equal(p, rsl_rl_ref.feistel_permutation(n, seed=1, iteration=3))


def test_gae_restatement_against_closed_form():
    """compute_returns on a single env without dones: returns_t - V_t = sum_k (gamma lam)^k delta_{t+k}."""
    T, g, l = 5, 0.99, 0.95
    s = rsl_rl_ref.RolloutStorage(1, T, [3], [None], [2])
    torch.manual_seed(0)
    s.rewards.copy_(torch.randn(T, 1, 1)); s.values.copy_(torch.randn(T, 1, 1))
    last = torch.randn(1, 1)
    s.compute_returns(last, g, l)
    v = torch.cat([s.values.flatten(), last.flatten()]).double()
    r = s.rewards.flatten().double()
    delta = r + g * v[1:] - v[:-1]
    for t in range(T):
        adv = sum((g * l) ** k * delta[t + k] for k in range(T - t))
        assert abs(float(s.returns[t]) - float(adv + v[t])) < 1e-5


def test_oracle_ppo_iteration_config_c1():
    """BASELINE.json configs[0] shape: 64 envs, one PPO iteration (24 steps, GAE, 5 x 4 minibatches), [128,64,32]."""
    torch.manual_seed(0)
    ac = rsl_rl_ref.ActorCritic(27, 27, 6, [128, 64, 32], [128, 64, 32])
    ppo = rsl_rl_ref.PPO(ac, num_learning_epochs=5, num_mini_batches=4, gamma=0.99, lam=0.95, entropy_coef=0.01,
                         schedule="adaptive", desired_kl=0.01)
    ppo.init_storage(64, 24, [27], [None], [6])
    p0 = torch.cat([p.detach().flatten() for p in ac.parameters()]).clone()
    for t in range(24):
        obs = torch.randn(64, 27)
        with torch.no_grad():
            ppo.act(obs, obs)
        ppo.process_env_step(torch.randn(64) * 0.1, torch.rand(64) < 0.05, {"time_outs": torch.zeros(64, dtype=torch.bool)})
    with pytest.raises(AssertionError, match="Rollout buffer overflow"):
        with torch.no_grad():
            ppo.act(obs, obs)
        ppo.process_env_step(torch.zeros(64), torch.zeros(64, dtype=torch.bool), {})
    with torch.no_grad():
        ppo.compute_returns(torch.randn(64, 27))
    vl, sl = ppo.update()
    p1 = torch.cat([p.detach().flatten() for p in ac.parameters()])
    assert np.isfinite(vl) and np.isfinite(sl) and (p1 - p0).abs().max() > 0 and len(ppo.last_kl) == 20
    assert abs(float(ppo.storage.advantages.mean())) < 1e-5 and abs(float(ppo.storage.advantages.std()) - 1) < 1e-4
