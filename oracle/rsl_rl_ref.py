"""Plain-torch restatement of rsl_rl v1.0.2 (ActorCritic / RolloutStorage / PPO / OnPolicyRunner).
ORACLE / TEST CODE -- the checker and the timed CPU baseline, never the product path.

**parity unpinned**: rsl_rl is a third-party dependency of the reference (README.md:34-36 pins tag v1.0.2;
call sites legged_gym/utils/task_registry.py:37-38,147,154, legged_gym/scripts/train.py:44, play.py:59) whose
source is NOT under /root/reference and is not installed here.  What follows restates the published algorithm
of that tag module by module; nothing in the reference pins its numbers.
    modules/actor_critic.py   -> ActorCritic
    storage/rollout_storage.py -> RolloutStorage
    algorithms/ppo.py         -> PPO
    runners/on_policy_runner.py -> OnPolicyRunner (learn loop only; TensorBoard logging left out)

Random draws: ``Normal.sample`` and ``torch.randperm`` are replaced by the counter-based contract of oracle/rng.py
(purposes POLICY and the Feistel minibatch permutation below) so that the CUDA path can be compared bit for bit
on what it samples; upstream draws from torch's global generator, which no implementation could replay.
"""
import math
from collections import deque

import numpy as np
import torch
import torch.nn as nn
from torch.distributions import Normal

from . import rng

MINIBATCH_PURPOSE = 7     # the Feistel permutation key is derived from (seed, iteration, 7)


def get_activation(name):
    return {"elu": nn.ELU(), "selu": nn.SELU(), "relu": nn.ReLU(), "lrelu": nn.LeakyReLU(), "tanh": nn.Tanh(),
            "sigmoid": nn.Sigmoid()}[name]


# ------------------------------------------------------------------------------------------------
def feistel_permutation(n, seed, iteration):
    """The minibatch shuffle contract shared with the kernels (stands in for torch.randperm at
    rollout_storage.py mini_batch_generator): a 4-round Feistel network over ceil-to-even log2(n) bits with
    cycle walking.  Returns perm[i] for i in range(n) as int64; a bijection of range(n)."""
    bits = max(2, int(n - 1).bit_length())
    bits += bits & 1
    half = bits // 2
    mask = np.uint64((1 << half) - 1)
    keys = [rng.rand_u32(seed, iteration, MINIBATCH_PURPOSE, 0, r).astype(np.uint64) for r in range(4)]
    x = np.arange(n, dtype=np.uint64)
    todo = np.ones(n, dtype=bool)
    first = True
    while todo.any():
        v = x[todo]
        l, r = v >> np.uint64(half), v & mask
        for k in keys:
            f = rng._fmix32((r + k) & np.uint64(0xFFFFFFFF)) & mask
            l, r = r, l ^ f
        v = (l << np.uint64(half)) | r
        x[todo] = v
        todo[todo] = v >= np.uint64(n)
        first = False
    return x.astype(np.int64)


# ------------------------------------------------------------------------------------------------
class ActorCritic(nn.Module):
    """modules/actor_critic.py (ctor args at pointfoot_rough_config.py:220-224)."""
    is_recurrent = False

    def __init__(self, num_actor_obs, num_critic_obs, num_actions, actor_hidden_dims=(256, 256, 256),
                 critic_hidden_dims=(256, 256, 256), activation="elu", init_noise_std=1.0, **kwargs):
        super().__init__()
        act = get_activation(activation)

        def mlp(inp, hidden, out):
            layers = [nn.Linear(inp, hidden[0]), act]
            for l in range(len(hidden)):
                if l == len(hidden) - 1:
                    layers.append(nn.Linear(hidden[l], out))
                else:
                    layers.append(nn.Linear(hidden[l], hidden[l + 1]))
                    layers.append(act)
            return nn.Sequential(*layers)
        self.actor = mlp(num_actor_obs, list(actor_hidden_dims), num_actions)
        self.critic = mlp(num_critic_obs, list(critic_hidden_dims), 1)
        self.std = nn.Parameter(init_noise_std * torch.ones(num_actions))
        self.distribution = None
        Normal.set_default_validate_args = False
        # sampling contract (not upstream): see module docstring
        self.sample_seed, self.sample_step, self.env_id_offset = 1, 0, 0

    def reset(self, dones=None):
        pass

    @property
    def action_mean(self):
        return self.distribution.mean

    @property
    def action_std(self):
        return self.distribution.stddev

    @property
    def entropy(self):
        return self.distribution.entropy().sum(dim=-1)

    def update_distribution(self, observations):
        mean = self.actor(observations)
        self.distribution = Normal(mean, mean * 0. + self.std)

    def act(self, observations, **kwargs):
        self.update_distribution(observations)
        n, a = self.distribution.mean.shape
        self.sample_step += 1
        z = rng.normal(self.sample_seed, self.sample_step, rng.POLICY,
                       (np.arange(n, dtype=np.int64) + self.env_id_offset)[:, None], np.arange(a)[None, :])
        with torch.no_grad():
            return self.distribution.mean + self.distribution.stddev * torch.from_numpy(z)

    def get_actions_log_prob(self, actions):
        return self.distribution.log_prob(actions).sum(dim=-1)

    def act_inference(self, observations):
        return self.actor(observations)

    def evaluate(self, critic_observations, **kwargs):
        return self.critic(critic_observations)


# ------------------------------------------------------------------------------------------------
class RolloutStorage:
    """storage/rollout_storage.py."""

    class Transition:
        def __init__(self):
            self.clear()

        def clear(self):
            self.observations = self.critic_observations = self.actions = self.rewards = self.dones = None
            self.values = self.actions_log_prob = self.action_mean = self.action_sigma = None

    def __init__(self, num_envs, num_transitions_per_env, obs_shape, privileged_obs_shape, actions_shape, device="cpu"):
        T, N = num_transitions_per_env, num_envs
        self.device = device
        z = lambda *s: torch.zeros(T, N, *s, device=device)
        self.observations = z(*obs_shape)
        self.privileged_observations = z(*privileged_obs_shape) if privileged_obs_shape[0] is not None else None
        self.rewards, self.actions = z(1), z(*actions_shape)
        self.dones = z(1).byte()
        self.actions_log_prob, self.values, self.returns, self.advantages = z(1), z(1), z(1), z(1)
        self.mu, self.sigma = z(*actions_shape), z(*actions_shape)
        self.num_transitions_per_env, self.num_envs = T, N
        self.step = 0

    def add_transitions(self, t):
        if self.step >= self.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        s = self.step
        self.observations[s].copy_(t.observations)
        if self.privileged_observations is not None:
            self.privileged_observations[s].copy_(t.critic_observations)
        self.actions[s].copy_(t.actions)
        self.rewards[s].copy_(t.rewards.view(-1, 1))
        self.dones[s].copy_(t.dones.view(-1, 1))
        self.values[s].copy_(t.values)
        self.actions_log_prob[s].copy_(t.actions_log_prob.view(-1, 1))
        self.mu[s].copy_(t.action_mean)
        self.sigma[s].copy_(t.action_sigma)
        self.step += 1

    def clear(self):
        self.step = 0

    def compute_returns(self, last_values, gamma, lam):
        advantage = 0
        for step in reversed(range(self.num_transitions_per_env)):
            next_values = last_values if step == self.num_transitions_per_env - 1 else self.values[step + 1]
            next_is_not_terminal = 1.0 - self.dones[step].float()
            delta = self.rewards[step] + next_is_not_terminal * gamma * next_values - self.values[step]
            advantage = delta + next_is_not_terminal * gamma * lam * advantage
            self.returns[step] = advantage + self.values[step]
        self.advantages = self.returns - self.values
        self.advantages = (self.advantages - self.advantages.mean()) / (self.advantages.std() + 1e-8)

    def mini_batch_generator(self, num_mini_batches, num_epochs=8, seed=1, iteration=0):
        batch_size = self.num_envs * self.num_transitions_per_env
        mini_batch_size = batch_size // num_mini_batches
        indices = torch.from_numpy(feistel_permutation(num_mini_batches * mini_batch_size, seed, iteration))
        flat = lambda x: x.flatten(0, 1)
        obs = flat(self.observations)
        cobs = flat(self.privileged_observations) if self.privileged_observations is not None else obs
        for epoch in range(num_epochs):
            for i in range(num_mini_batches):
                b = indices[i * mini_batch_size:(i + 1) * mini_batch_size]
                yield (obs[b], cobs[b], flat(self.actions)[b], flat(self.values)[b], flat(self.advantages)[b],
                       flat(self.returns)[b], flat(self.actions_log_prob)[b], flat(self.mu)[b], flat(self.sigma)[b],
                       (None, None), None)


# ------------------------------------------------------------------------------------------------
class PPO:
    """algorithms/ppo.py (hyper-parameters at pointfoot_rough_config.py:226-243)."""

    def __init__(self, actor_critic, num_learning_epochs=1, num_mini_batches=1, clip_param=0.2, gamma=0.998, lam=0.95,
                 value_loss_coef=1.0, entropy_coef=0.0, learning_rate=1e-3, max_grad_norm=1.0,
                 use_clipped_value_loss=True, schedule="fixed", desired_kl=0.01, device="cpu"):
        self.device = device
        self.desired_kl, self.schedule, self.learning_rate = desired_kl, schedule, learning_rate
        self.actor_critic = actor_critic
        self.storage = None
        self.optimizer = torch.optim.Adam(self.actor_critic.parameters(), lr=learning_rate)
        self.transition = RolloutStorage.Transition()
        self.clip_param, self.num_learning_epochs, self.num_mini_batches = clip_param, num_learning_epochs, num_mini_batches
        self.value_loss_coef, self.entropy_coef, self.gamma, self.lam = value_loss_coef, entropy_coef, gamma, lam
        self.max_grad_norm, self.use_clipped_value_loss = max_grad_norm, use_clipped_value_loss
        self.seed, self.iteration = 1, 0
        self.last_kl = []                 # checker hook: kl_mean per minibatch
        self.last_grad = None             # checker hook: flat gradient of the last minibatch (before clipping)

    def init_storage(self, num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape, action_shape):
        self.storage = RolloutStorage(num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape,
                                      action_shape, self.device)

    def test_mode(self):
        self.actor_critic.eval()

    def train_mode(self):
        self.actor_critic.train()

    def act(self, obs, critic_obs):
        t = self.transition
        t.actions = self.actor_critic.act(obs).detach()
        t.values = self.actor_critic.evaluate(critic_obs).detach()
        t.actions_log_prob = self.actor_critic.get_actions_log_prob(t.actions).detach()
        t.action_mean = self.actor_critic.action_mean.detach()
        t.action_sigma = self.actor_critic.action_std.detach()
        t.observations, t.critic_observations = obs, critic_obs
        return t.actions

    def process_env_step(self, rewards, dones, infos):
        t = self.transition
        t.rewards = rewards.clone()
        t.dones = dones
        if "time_outs" in infos:
            t.rewards += self.gamma * torch.squeeze(t.values * infos["time_outs"].unsqueeze(1).to(self.device), 1)
        self.storage.add_transitions(t)
        t.clear()
        self.actor_critic.reset(dones)

    def compute_returns(self, last_critic_obs):
        last_values = self.actor_critic.evaluate(last_critic_obs).detach()
        self.storage.compute_returns(last_values, self.gamma, self.lam)

    def update(self):
        mean_value_loss = mean_surrogate_loss = 0
        self.last_kl = []
        gen = self.storage.mini_batch_generator(self.num_mini_batches, self.num_learning_epochs, self.seed, self.iteration)
        for (obs_b, cobs_b, act_b, tv_b, adv_b, ret_b, old_lp_b, old_mu_b, old_sigma_b, _, _) in gen:
            ac = self.actor_critic
            ac.update_distribution(obs_b)             # upstream calls act(); the sample it draws is discarded
            lp_b = ac.get_actions_log_prob(act_b)
            value_b = ac.evaluate(cobs_b)
            mu_b, sigma_b, entropy_b = ac.action_mean, ac.action_std, ac.entropy
            if self.desired_kl is not None and self.schedule == "adaptive":
                with torch.inference_mode():
                    kl = torch.sum(torch.log(sigma_b / old_sigma_b + 1.e-5)
                                   + (torch.square(old_sigma_b) + torch.square(old_mu_b - mu_b))
                                   / (2.0 * torch.square(sigma_b)) - 0.5, axis=-1)
                    kl_mean = torch.mean(kl)
                    self.last_kl.append(float(kl_mean))
                    if kl_mean > self.desired_kl * 2.0:
                        self.learning_rate = max(1e-5, self.learning_rate / 1.5)
                    elif kl_mean < self.desired_kl / 2.0 and kl_mean > 0.0:
                        self.learning_rate = min(1e-2, self.learning_rate * 1.5)
                    for g in self.optimizer.param_groups:
                        g["lr"] = self.learning_rate
            ratio = torch.exp(lp_b - torch.squeeze(old_lp_b))
            surrogate = -torch.squeeze(adv_b) * ratio
            surrogate_clipped = -torch.squeeze(adv_b) * torch.clamp(ratio, 1.0 - self.clip_param, 1.0 + self.clip_param)
            surrogate_loss = torch.max(surrogate, surrogate_clipped).mean()
            if self.use_clipped_value_loss:
                value_clipped = tv_b + (value_b - tv_b).clamp(-self.clip_param, self.clip_param)
                value_loss = torch.max((value_b - ret_b).pow(2), (value_clipped - ret_b).pow(2)).mean()
            else:
                value_loss = (ret_b - value_b).pow(2).mean()
            loss = surrogate_loss + self.value_loss_coef * value_loss - self.entropy_coef * entropy_b.mean()
            self.optimizer.zero_grad()
            loss.backward()
            self.last_grad = torch.cat([p.grad.flatten() for p in ac.parameters()]).clone()
            nn.utils.clip_grad_norm_(ac.parameters(), self.max_grad_norm)
            self.optimizer.step()
            mean_value_loss += value_loss.item()
            mean_surrogate_loss += surrogate_loss.item()
        n = self.num_learning_epochs * self.num_mini_batches
        self.storage.clear()
        self.iteration += 1
        return mean_value_loss / n, mean_surrogate_loss / n


# ------------------------------------------------------------------------------------------------
class OnPolicyRunner:
    """runners/on_policy_runner.py: the learn loop (logging / checkpoints left to the caller)."""

    def __init__(self, env, train_cfg, log_dir=None, device="cpu"):
        self.cfg, self.alg_cfg, self.policy_cfg = train_cfg["runner"], train_cfg["algorithm"], train_cfg["policy"]
        self.device, self.env = device, env
        num_critic_obs = env.num_privileged_obs if env.num_privileged_obs is not None else env.num_obs
        ac = ActorCritic(env.num_obs, num_critic_obs, env.num_actions, **self.policy_cfg).to(device)
        self.alg = PPO(ac, device=device, **self.alg_cfg)
        self.num_steps_per_env, self.save_interval = self.cfg["num_steps_per_env"], self.cfg["save_interval"]
        self.alg.init_storage(env.num_envs, self.num_steps_per_env, [env.num_obs], [env.num_privileged_obs],
                              [env.num_actions])
        self.log_dir, self.current_learning_iteration = log_dir, 0
        self.rewbuffer, self.lenbuffer = deque(maxlen=100), deque(maxlen=100)
        self.env.reset()

    def learn(self, num_learning_iterations, init_at_random_ep_len=False):
        env = self.env
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        critic_obs = priv if priv is not None else obs
        self.alg.actor_critic.train()
        cur_rew = torch.zeros(env.num_envs)
        cur_len = torch.zeros(env.num_envs)
        out = []
        for it in range(self.current_learning_iteration, self.current_learning_iteration + num_learning_iterations):
            with torch.inference_mode():
                for _ in range(self.num_steps_per_env):
                    actions = self.alg.act(obs, critic_obs)
                    obs, priv, rewards, dones, infos = env.step(actions)
                    critic_obs = priv if priv is not None else obs
                    self.alg.process_env_step(rewards, dones, infos)
                    cur_rew += rewards
                    cur_len += 1
                    new_ids = (dones > 0).nonzero(as_tuple=False)
                    self.rewbuffer.extend(cur_rew[new_ids][:, 0].numpy().tolist())
                    self.lenbuffer.extend(cur_len[new_ids][:, 0].numpy().tolist())
                    cur_rew[new_ids] = 0
                    cur_len[new_ids] = 0
                self.alg.compute_returns(critic_obs)
            out.append(self.alg.update())
        self.current_learning_iteration += num_learning_iterations
        return out

    def get_inference_policy(self, device=None):
        self.alg.actor_critic.eval()
        return self.alg.actor_critic.act_inference
