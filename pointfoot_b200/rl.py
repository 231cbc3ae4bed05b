"""rsl_rl v1.0.2-compatible ``ActorCritic`` / ``RolloutStorage`` / ``PPO`` / ``OnPolicyRunner`` whose arithmetic is
launches into libpointfoot_b200.so (include/pointfoot_b200.h, "rsl_rl side").

Same class names, constructor arguments, methods and checkpoint format as the upstream package the reference
imports (legged_gym/utils/task_registry.py:37-38,147,154; legged_gym/scripts/train.py:44; play.py:59,64), so that
``task_registry.make_alg_runner`` and ``runner.learn()`` work unchanged.  There is no torch arithmetic on the path:
torch only owns the device memory (parameters, rollout storage) and the streams, and -- on more than one GPU --
``torch.distributed`` (NCCL) carries the gradient / statistic all-reduces.  No CPU fallback.

Deliberate differences from upstream, all about randomness (upstream draws from torch's global generator, which
cannot be replayed): ``Normal.sample`` and the minibatch ``randperm`` follow the counter-based contract of
csrc/pf_rng.cuh (purposes POLICY and 7).
"""
import ctypes as C
import os
import statistics
import time
from collections import deque

import torch

from . import capi, shard


def _ptr(t):
    return C.c_void_p(None if t is None else t.data_ptr())


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _checked(t, device, cols, what, rows=None, dtype=torch.float32):
    """The kernels are handed raw pointers: refuse anything that is not the dense tensor they index.  Returns ``t``
    (made contiguous if it is a strided view of the right dtype / device)."""
    if not torch.is_tensor(t):
        raise TypeError(f"{what}: expected a torch tensor, got {type(t).__name__}")
    if t.device != device:
        raise RuntimeError(f"{what}: tensor is on {t.device}, the policy lives on {device}")
    if t.dtype != dtype:
        raise RuntimeError(f"{what}: dtype {t.dtype}, expected {dtype}")
    if cols is not None and (t.dim() < 1 or t.numel() != t.shape[0] * cols):
        raise RuntimeError(f"{what}: shape {tuple(t.shape)} does not have {cols} columns")
    if rows is not None and t.shape[0] != rows:
        raise RuntimeError(f"{what}: {t.shape[0]} rows, expected {rows}")
    return t if t.is_contiguous() else t.contiguous()


class ActorCritic:
    """modules/actor_critic.py.  Parameters live in ONE flat fp32 device buffer in ``parameters()`` order
    (std, actor.*, critic.*); ``state_dict()`` exposes the upstream key names as views of it."""
    is_recurrent = False

    def __init__(self, num_actor_obs, num_critic_obs, num_actions, actor_hidden_dims=(256, 256, 256),
                 critic_hidden_dims=(256, 256, 256), activation="elu", init_noise_std=1.0, device="cuda:0", **kwargs):
        if activation != "elu":
            raise NotImplementedError("pointfoot_b200: only the 'elu' activation of the pointfoot configs is implemented")
        self.device = torch.device(device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("pointfoot_b200.rl needs a CUDA device (no CPU fallback)")
        self.num_actor_obs, self.num_critic_obs, self.num_actions = num_actor_obs, num_critic_obs, num_actions
        self.actor_dims = [num_actor_obs] + list(actor_hidden_dims) + [num_actions]
        self.critic_dims = [num_critic_obs] + list(critic_hidden_dims) + [1]
        self.init_noise_std = init_noise_std
        self.training = True
        self.flat = None          # allocated by bind()
        self._views = None
        self._ctx = None

    # -- layout ------------------------------------------------------------------------------------
    def layout(self):
        """[(state_dict key, offset, shape)] in ActorCritic.parameters() order."""
        out, off = [("std", 0, (self.num_actions,))], self.num_actions
        for name, dims in (("actor", self.actor_dims), ("critic", self.critic_dims)):
            for l in range(len(dims) - 1):
                out.append((f"{name}.{2 * l}.weight", off, (dims[l + 1], dims[l]))); off += dims[l] * dims[l + 1]
                out.append((f"{name}.{2 * l}.bias", off, (dims[l + 1],))); off += dims[l + 1]
        return out, off

    def bind(self, lib, ctx):
        """Attach to a pf_rl_ctx: allocate the flat buffer and initialise it like torch.nn.Linear does."""
        self._lib, self._ctx = lib, ctx
        lay, n = self.layout()
        assert n == lib.pf_rl_param_count(ctx), "parameter layout mismatch with the library"
        self.flat = torch.zeros(n, device=self.device)
        self._views = {k: self.flat[o:o + int(torch.Size(s).numel())].view(s) for k, o, s in lay}
        self.reset_parameters()

    def reset_parameters(self, generator=None):
        """nn.Linear default init (kaiming_uniform(a=sqrt(5)) = U(+-1/sqrt(fan_in)) for weight and bias), std = init."""
        for k, v in self._views.items():
            if k == "std":
                v.fill_(self.init_noise_std)
                continue
            name, idx, kind = k.split(".")
            dims = self.actor_dims if name == "actor" else self.critic_dims
            bound = 1.0 / (dims[int(idx) // 2] ** 0.5)
            v.copy_(((torch.rand(v.shape, generator=generator) * 2 - 1) * bound).to(self.device))
        self.params_changed()

    def params_changed(self):
        capi.check(self._lib.pf_rl_params_changed(self._ctx, _ptr(self.flat), _stream(self.device)), "pf_rl_params_changed")

    # -- nn.Module surface the callers use ---------------------------------------------------------
    @property
    def std(self):
        return self._views["std"]

    def parameters(self):
        return [self._views[k] for k, _, _ in self.layout()[0]]

    def state_dict(self):
        return {k: v for k, v in self._views.items()}

    def load_state_dict(self, sd, strict=True):
        for k, v in self._views.items():
            if k in sd:
                v.copy_(sd[k].to(self.device, torch.float).view(v.shape))
            elif strict:
                raise KeyError(k)
        self.params_changed()

    def train(self, mode=True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)

    def to(self, device):
        return self

    def reset(self, dones=None):
        pass

    # -- inference ---------------------------------------------------------------------------------
    def act_inference(self, observations):
        """actor(observations) for ANY number of rows (upstream: a plain nn.Sequential call)"""
        obs = _checked(observations, self.flat.device, self.num_actor_obs, "act_inference(observations)")
        out = torch.empty(obs.shape[0], self.num_actions, device=self.device)
        capi.check(self._lib.pf_rl_act_inference(self._ctx, _ptr(self.flat), _ptr(obs), C.c_int32(obs.shape[0]), _ptr(out),
                                                 _stream(self.device)), "pf_rl_act_inference")
        return out

    def evaluate(self, critic_observations, **kwargs):
        cobs = _checked(critic_observations, self.flat.device, self.num_critic_obs, "evaluate(critic_observations)")
        out = torch.empty(cobs.shape[0], 1, device=self.device)
        capi.check(self._lib.pf_rl_evaluate(self._ctx, _ptr(self.flat), _ptr(cobs), C.c_int32(cobs.shape[0]), _ptr(out),
                                            _stream(self.device)), "pf_rl_evaluate")
        return out


class RolloutStorage:
    """storage/rollout_storage.py: the [T, N, .] tensors (torch owns the memory; the kernels fill them)."""

    def __init__(self, num_envs, num_transitions_per_env, obs_shape, privileged_obs_shape, actions_shape, device="cuda:0"):
        T, N = num_transitions_per_env, num_envs
        self.device = device
        z = lambda *s, dt=torch.float: torch.zeros(T, N, *s, device=device, dtype=dt)
        self.observations = z(*obs_shape)
        self.privileged_observations = z(*privileged_obs_shape) if privileged_obs_shape[0] is not None else None
        self.rewards, self.actions = z(1), z(*actions_shape)
        self.dones = z(1, dt=torch.uint8)
        self.actions_log_prob, self.values, self.returns, self.advantages = z(1), z(1), z(1), z(1)
        self.mu, self.sigma = z(*actions_shape), z(*actions_shape)
        self.num_transitions_per_env, self.num_envs = T, N
        self.step = 0
        self.c = capi.PfRlStorage()
        for n in capi.RL_STORAGE_FIELDS:
            t = getattr(self, n)
            setattr(self.c, n, None if t is None else t.data_ptr())

    def clear(self):
        self.step = 0




class PPO:
    """algorithms/ppo.py: act / process_env_step / compute_returns / update, each a handful of kernel launches."""

    def __init__(self, actor_critic, num_learning_epochs=1, num_mini_batches=1, clip_param=0.2, gamma=0.998, lam=0.95,
                 value_loss_coef=1.0, entropy_coef=0.0, learning_rate=1e-3, max_grad_norm=1.0,
                 use_clipped_value_loss=True, schedule="fixed", desired_kl=0.01, device="cuda:0", seed=1,
                 precision=None, env_id_offset=0):
        self.device = torch.device(device)
        self.actor_critic = actor_critic
        self.desired_kl, self.schedule = desired_kl, schedule
        self.clip_param, self.num_learning_epochs, self.num_mini_batches = clip_param, num_learning_epochs, num_mini_batches
        self.value_loss_coef, self.entropy_coef, self.gamma, self.lam = value_loss_coef, entropy_coef, gamma, lam
        self.max_grad_norm, self.use_clipped_value_loss = max_grad_norm, use_clipped_value_loss
        self._lr0 = learning_rate
        self.seed, self.iteration, self.policy_step, self.env_id_offset = seed, 0, 0, env_id_offset
        precision = precision or os.environ.get("PF_RL_PRECISION", "fp32")
        self.precision = {"fp32": capi.RL_FP32, "bf16": capi.RL_BF16_TC}[precision]
        self.storage, self._ctx, self._t = None, None, 0
        self._update_graph, self._graph_updates = None, os.environ.get("PF_RL_UPDATE_GRAPH", "1") != "0"
        self.launches_per_optimizer_step = None          # measured (library launch counter) by the first update()
        self._graph_dist = os.environ.get("PF_RL_UPDATE_GRAPH_DIST", "1") != "0"
        self._lib = capi.load_rl()

    # learning_rate mirrors the device scalar the adaptive schedule updates (one sync when read)
    @property
    def learning_rate(self):
        return float(self._lr.item()) if self._ctx else self._lr0

    @learning_rate.setter
    def learning_rate(self, v):
        if self._ctx:
            self._lr.fill_(v)
        else:
            self._lr0 = v

    def init_storage(self, num_envs, num_transitions_per_env, actor_obs_shape, critic_obs_shape, action_shape):
        ac = self.actor_critic
        P = capi.PfRlParams()
        P.abi_version, P.num_envs, P.env_id_offset, P.num_steps = capi.PF_ABI_VERSION, num_envs, self.env_id_offset, num_transitions_per_env
        P.num_obs, P.num_critic_obs, P.num_actions = ac.num_actor_obs, ac.num_critic_obs, ac.num_actions
        P.actor_layers, P.critic_layers = len(ac.actor_dims) - 1, len(ac.critic_dims) - 1
        for i, d in enumerate(ac.actor_dims):
            P.actor_dims[i] = d
        for i, d in enumerate(ac.critic_dims):
            P.critic_dims[i] = d
        P.seed = self.seed
        P.gamma, P.lam, P.clip_param = self.gamma, self.lam, self.clip_param
        P.value_loss_coef, P.entropy_coef, P.max_grad_norm = self.value_loss_coef, self.entropy_coef, self.max_grad_norm
        P.use_clipped_value_loss = int(self.use_clipped_value_loss)
        P.adaptive_schedule = int(self.desired_kl is not None and self.schedule == "adaptive")
        P.num_mini_batches, P.num_learning_epochs, P.precision = self.num_mini_batches, self.num_learning_epochs, self.precision
        P.desired_kl = float(self.desired_kl or 0.0)
        P.adam_beta1, P.adam_beta2, P.adam_eps = 0.9, 0.999, 1e-8            # torch.optim.Adam defaults
        self._P = P
        ctx = C.c_void_p()
        with torch.cuda.device(self.device):
            capi.check(self._lib.pf_rl_create(C.byref(P), C.byref(ctx)), "pf_rl_create")
        self._ctx = ctx
        # which kernels serve act / inference / evaluate (include/pointfoot_b200.h: pf_rl_inference_path)
        self.inference_path = {0: "fp32", 1: "tcgen05", 2: "warp_mma"}[self._lib.pf_rl_inference_path(ctx)]
        ac.bind(self._lib, ctx)
        d = shard.dist_or_none()
        if d:                                      # every shard starts from rank 0's initial parameters
            d.broadcast(ac.flat, 0)
            ac.params_changed()
        crit = critic_obs_shape if critic_obs_shape[0] is not None else [None]
        self.storage = RolloutStorage(num_envs, num_transitions_per_env, actor_obs_shape, crit, action_shape, self.device)
        n, dev = ac.flat.numel(), self.device
        self.grads, self.adam_m, self.adam_v = (torch.zeros(n, device=dev) for _ in range(3))
        self._lr = torch.full((1,), self._lr0, dtype=torch.float64, device=dev)
        self.adam_step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.loss_stats = torch.zeros(8, dtype=torch.float64, device=dev)
        self._gae_stats = torch.zeros(3, dtype=torch.float64, device=dev)
        self._last_values = torch.zeros(num_envs, device=dev)
        self._actions = torch.zeros(num_envs, ac.num_actions, device=dev)
        self._loss_acc = torch.zeros(8, dtype=torch.float64, device=dev)
        L = capi.PfRlLearner()
        for k, t in (("params", ac.flat), ("grads", self.grads), ("adam_m", self.adam_m), ("adam_v", self.adam_v),
                     ("lr", self._lr), ("adam_step", self.adam_step), ("loss_stats", self.loss_stats)):
            setattr(L, k, t.data_ptr())
        self._L = L

    def test_mode(self):
        self.actor_critic.eval()

    def train_mode(self):
        self.actor_critic.train()

    def _st(self):
        return _stream(self.device)

    def act(self, obs, critic_obs):
        """ppo.py act + the storage copies of add_transitions; returns the sampled actions [N, A]."""
        if self.storage.step >= self.storage.num_transitions_per_env:
            raise AssertionError("Rollout buffer overflow")
        ac, N = self.actor_critic, self.storage.num_envs
        obs = _checked(obs, ac.flat.device, ac.num_actor_obs, "PPO.act(obs)", N)
        critic_obs = _checked(critic_obs, ac.flat.device, ac.num_critic_obs, "PPO.act(critic_obs)", N)
        self.policy_step += 1
        capi.check(self._lib.pf_rl_act(self._ctx, _ptr(self.actor_critic.flat), _ptr(obs), _ptr(critic_obs),
                                       C.c_int32(self.storage.step), C.byref(self.storage.c), C.c_int64(self.policy_step),
                                       _ptr(self._actions), None, None, self._st()), "pf_rl_act")
        return self._actions

    def process_env_step(self, rewards, dones, infos):
        to = infos.get("time_outs") if isinstance(infos, dict) else None
        dev, N = self.actor_critic.flat.device, self.storage.num_envs
        rewards = _checked(rewards, dev, 1, "process_env_step(rewards)", N)
        if dones.dtype not in (torch.bool, torch.uint8):
            raise RuntimeError(f"process_env_step(dones): dtype {dones.dtype}, expected torch.bool / uint8")
        dones = _checked(dones, dev, 1, "process_env_step(dones)", N, dones.dtype)
        if to is not None:
            to = _checked(to, dev, 1, "process_env_step(time_outs)", N, torch.bool if to.dtype == torch.bool else torch.uint8)
        capi.check(self._lib.pf_rl_process_env_step(self._ctx, C.byref(self.storage.c), C.c_int32(self.storage.step),
                                                    _ptr(rewards), _ptr(dones), _ptr(to), self._st()), "pf_rl_process_env_step")
        self.storage.step += 1

    def compute_returns(self, last_critic_obs):
        lib, st = self._lib, self._st()
        ac, N = self.actor_critic, self.storage.num_envs
        last_critic_obs = _checked(last_critic_obs, ac.flat.device, ac.num_critic_obs, "compute_returns(last_critic_obs)", N)
        capi.check(lib.pf_rl_evaluate(self._ctx, _ptr(ac.flat), _ptr(last_critic_obs), C.c_int32(N), _ptr(self._last_values), st),
                   "pf_rl_evaluate")
        capi.check(lib.pf_rl_gae(self._ctx, C.byref(self.storage.c), _ptr(self._last_values), _ptr(self._gae_stats), st), "pf_rl_gae")
        shard.allreduce_sum_(self._gae_stats)              # advantage mean / std span all shards
        capi.check(lib.pf_rl_normalize_advantages(self._ctx, C.byref(self.storage.c), _ptr(self._gae_stats), st),
                   "pf_rl_normalize_advantages")

    def minibatch_grad(self, mb):
        capi.check(self._lib.pf_rl_minibatch_grad(self._ctx, C.byref(self.storage.c), C.byref(self._L), C.c_int32(mb), self._st()),
                   "pf_rl_minibatch_grad")
        if shard.dist_or_none():   # one flat buffer -> ONE all-reduce per optimizer step (1.68 MB rough / 112 KB flat)
            shard.allreduce_mean_(self.grads)
            shard.allreduce_sum_(self.loss_stats)

    def _update_steps(self):
        """num_learning_epochs x num_mini_batches optimizer steps (ppo.py update(), the body of its generator loop)"""
        lib = self._lib
        for _ in range(self.num_learning_epochs):
            for mb in range(self.num_mini_batches):
                self.minibatch_grad(mb)
                capi.check(lib.pf_rl_adam_step(self._ctx, C.byref(self._L), self._st()), "pf_rl_adam_step")
                self._loss_acc += self.loss_stats

    def update(self):
        capi.check(self._lib.pf_rl_make_permutation(self._ctx, C.c_int64(self.iteration), self._st()), "pf_rl_make_permutation")
        self._loss_acc.zero_()
        # All optimizer steps of one update are a fixed sequence of ~2,000 small launches on fixed buffers (learning
        # rate, Adam state and the permutation live on the device): captured once, replayed as ONE graph launch.
        # With more than one rank the NCCL all-reduces stay outside a graph, so the steps are launched one by one.
        n_steps = self.num_learning_epochs * self.num_mini_batches
        # More than one rank: the NCCL all-reduces of the flat gradient and of loss_stats are captured with the kernels (the
        # communicator exists since init_storage's broadcast), so the multi-GPU update is ONE graph launch as well.
        # PF_RL_UPDATE_GRAPH_DIST=0 (or a failed capture) falls back to launching the steps one by one.
        use_graph = self._graph_updates and (shard.dist_or_none() is None or self._graph_dist)
        if use_graph and self._update_graph is None:
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            n0 = self._lib.pf_debug_launch_count()
            try:
                with torch.cuda.graph(g):
                    self._update_steps()
                self._update_graph = g
                self.launches_per_optimizer_step = (self._lib.pf_debug_launch_count() - n0) / n_steps
            except Exception as ex:                    # capture of the collectives refused: eager from here on
                if shard.dist_or_none() is None:
                    raise
                print(f"pointfoot_b200: PPO.update graph capture with collectives failed ({type(ex).__name__}: {ex}); "
                      "launching the update steps eagerly")
                self._graph_dist, use_graph = False, False
                torch.cuda.synchronize(self.device)
                self._loss_acc.zero_()
        if use_graph:
            self._update_graph.replay()
            # the replayed graph ends in the kernels that re-pack the inference weights: the next act must not read them early
            self._lib.pf_rl_weights_written(self._ctx)
        else:
            n0 = self._lib.pf_debug_launch_count()
            self._update_steps()
            self.launches_per_optimizer_step = (self._lib.pf_debug_launch_count() - n0) / n_steps
        acc = self._loss_acc.tolist()                      # the one host sync of update(), like upstream's .item()
        self.storage.clear()
        self.iteration += 1
        cnt = max(acc[4], 1.0)
        return acc[1] / cnt, acc[0] / cnt

    def close(self):
        # the captured update holds this context's buffers and, on more than one rank, NCCL kernels: it has to go BEFORE the
        # context and before torch.distributed.destroy_process_group() (a live graph with captured collectives makes the
        # communicator's teardown wait forever)
        if self._update_graph is not None:
            torch.cuda.synchronize(self.device)
            self._update_graph = None
        if self._ctx:
            self._lib.pf_rl_destroy(self._ctx)
            self._ctx = None

    # -- torch.optim.Adam-compatible state (checkpoint interchange with the reference stack) -------
    def optimizer_state_dict(self):
        lay, _ = self.actor_critic.layout()
        state = {}
        for i, (k, o, s) in enumerate(lay):
            n = int(torch.Size(s).numel())
            state[i] = {"step": self.adam_step.to(torch.float32).reshape(()).cpu(),
                        "exp_avg": self.adam_m[o:o + n].view(s).clone(), "exp_avg_sq": self.adam_v[o:o + n].view(s).clone()}
        group = {"lr": self.learning_rate, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0, "amsgrad": False,
                 "params": list(range(len(lay)))}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, sd):
        lay, _ = self.actor_critic.layout()
        for i, (k, o, s) in enumerate(lay):
            st = sd["state"].get(i)
            if st is None:
                continue
            n = int(torch.Size(s).numel())
            self.adam_m[o:o + n].copy_(st["exp_avg"].to(self.device).reshape(-1))
            self.adam_v[o:o + n].copy_(st["exp_avg_sq"].to(self.device).reshape(-1))
            self.adam_step.fill_(int(float(st["step"])))
        if sd.get("param_groups"):
            self.learning_rate = sd["param_groups"][0]["lr"]


class _OptimizerView:
    """``runner.alg.optimizer`` as far as save()/load() use it."""
    def __init__(self, alg):
        self.alg = alg

    def state_dict(self):
        return self.alg.optimizer_state_dict()

    def load_state_dict(self, sd):
        self.alg.load_optimizer_state_dict(sd)


class OnPolicyRunner:
    """runners/on_policy_runner.py: same constructor, learn / save / load / get_inference_policy, same checkpoint
    dict (model_state_dict, optimizer_state_dict, iter, infos)."""

    def __init__(self, env, train_cfg, log_dir=None, device="cuda:0"):
        self.cfg, self.alg_cfg, self.policy_cfg = train_cfg["runner"], train_cfg["algorithm"], train_cfg["policy"]
        self.device, self.env = device, env
        num_critic_obs = env.num_privileged_obs if env.num_privileged_obs is not None else env.num_obs
        ac = ActorCritic(env.num_obs, num_critic_obs, env.num_actions, device=device, **self.policy_cfg)
        extra = {"seed": train_cfg.get("seed", 1), "env_id_offset": getattr(env, "env_id_offset", 0)}
        self.alg = PPO(ac, device=device, **self.alg_cfg, **extra)
        self.num_steps_per_env, self.save_interval = self.cfg["num_steps_per_env"], self.cfg["save_interval"]
        self.alg.init_storage(env.num_envs, self.num_steps_per_env, [env.num_obs], [env.num_privileged_obs], [env.num_actions])
        self.alg.optimizer = _OptimizerView(self.alg)
        self.log_dir, self.writer = log_dir, None
        self.tot_timesteps, self.tot_time, self.current_learning_iteration = 0, 0, 0
        self.last_losses = None
        self.env.reset()

    def learn(self, num_learning_iterations, init_at_random_ep_len=False):
        env, alg, dev = self.env, self.alg, self.device
        if self.log_dir is not None and self.writer is None:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self.writer = SummaryWriter(log_dir=self.log_dir, flush_secs=10)
            except Exception:                      # tensorboard is optional here; console logging still works
                self.writer = None
        if init_at_random_ep_len:
            env.episode_length_buf.copy_(torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length)))
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        critic_obs = priv if priv is not None else obs
        alg.actor_critic.train()
        ep_infos = []
        rewbuffer, lenbuffer = deque(maxlen=100), deque(maxlen=100)
        cur_reward_sum = torch.zeros(env.num_envs, dtype=torch.float, device=dev)
        cur_episode_length = torch.zeros(env.num_envs, dtype=torch.float, device=dev)
        tot_iter = self.current_learning_iteration + num_learning_iterations
        sharded = shard.dist_or_none() is not None
        for it in range(self.current_learning_iteration, tot_iter):
            start = time.time()
            for _ in range(self.num_steps_per_env):
                actions = alg.act(obs, critic_obs)
                obs, priv, rewards, dones, infos = env.step(actions)
                critic_obs = priv if priv is not None else obs
                alg.process_env_step(rewards, dones, infos)
                if sharded and "episode" in infos:     # every rank: the merge after the rollout is a collective
                    ep = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in infos["episode"].items()}
                    ep["_count"] = env._reset_count.clone() if hasattr(env, "_reset_count") else (dones > 0).sum()
                    ep_infos.append(ep)
                if self.log_dir is not None:
                    if "episode" in infos and not sharded:
                        ep_infos.append({k: (v.clone() if torch.is_tensor(v) else v) for k, v in infos["episode"].items()})
                    cur_reward_sum += rewards
                    cur_episode_length += 1
                    new_ids = (dones > 0).nonzero(as_tuple=False)
                    rewbuffer.extend(cur_reward_sum[new_ids][:, 0].cpu().numpy().tolist())
                    lenbuffer.extend(cur_episode_length[new_ids][:, 0].cpu().numpy().tolist())
                    cur_reward_sum[new_ids] = 0
                    cur_episode_length[new_ids] = 0
            alg.compute_returns(critic_obs)
            if sharded:                                # Episode/* of the whole job, not of this rank's shard
                ep_infos = [shard.allreduce_episode_infos(ep_infos, env.num_envs)] if ep_infos else []
            torch.cuda.synchronize(dev)
            stop = time.time()
            collection_time, start = stop - start, stop
            mean_value_loss, mean_surrogate_loss = alg.update()
            stop = time.time()
            learn_time = stop - start
            self.last_losses = (mean_value_loss, mean_surrogate_loss)
            self.last_times = (collection_time, learn_time)          # what runner.log() calls Perf/collection time, Perf/learning_time
            if self.log_dir is not None:
                self.log(locals())
                if it % self.save_interval == 0:
                    self.save(os.path.join(self.log_dir, "model_{}.pt".format(it)))
            ep_infos = []
        self.current_learning_iteration += num_learning_iterations
        if self.log_dir is not None:
            self.save(os.path.join(self.log_dir, "model_{}.pt".format(self.current_learning_iteration)))

    def log(self, locs, width=80, pad=35):
        n_steps = self.num_steps_per_env * self.env.num_envs
        self.tot_timesteps += n_steps
        iteration_time = locs["collection_time"] + locs["learn_time"]
        self.tot_time += iteration_time
        fps = int(n_steps / iteration_time)
        scalars = {"Loss/value_function": locs["mean_value_loss"], "Loss/surrogate": locs["mean_surrogate_loss"],
                   "Loss/learning_rate": self.alg.learning_rate, "Policy/mean_noise_std": float(self.alg.actor_critic.std.mean()),
                   "Perf/total_fps": fps, "Perf/collection time": locs["collection_time"], "Perf/learning_time": locs["learn_time"]}
        if locs["ep_infos"]:
            for key in locs["ep_infos"][0]:
                vals = [torch.as_tensor(ep[key], dtype=torch.float).reshape(-1).cpu() for ep in locs["ep_infos"]]
                scalars["Episode/" + key] = float(torch.cat(vals).mean())
        if len(locs["rewbuffer"]) > 0:
            scalars["Train/mean_reward"] = statistics.mean(locs["rewbuffer"])
            scalars["Train/mean_episode_length"] = statistics.mean(locs["lenbuffer"])
        if self.writer is not None:
            for k, v in scalars.items():
                self.writer.add_scalar(k, v, locs["it"])
        head = f" Learning iteration {locs['it']}/{locs['tot_iter']} "
        lines = ["#" * width, head.center(width, " "), ""]
        lines.append(f"{'Computation:':>{pad}} {fps:.0f} steps/s (collection: {locs['collection_time']:.3f}s, "
                     f"learning {locs['learn_time']:.3f}s)")
        for k, v in scalars.items():
            if k.startswith(("Loss/", "Policy/", "Train/", "Episode/")):
                lines.append(f"{k + ':':>{pad}} {v:.4f}")
        print("\n".join(lines))

    def save(self, path, infos=None):
        torch.save({"model_state_dict": {k: v.clone() for k, v in self.alg.actor_critic.state_dict().items()},
                    "optimizer_state_dict": self.alg.optimizer.state_dict(),
                    "iter": self.current_learning_iteration, "infos": infos}, path)

    def load(self, path, load_optimizer=True):
        d = torch.load(path, map_location="cpu", weights_only=False)
        self.alg.actor_critic.load_state_dict(d["model_state_dict"])
        if load_optimizer:
            self.alg.optimizer.load_state_dict(d["optimizer_state_dict"])
        self.current_learning_iteration = d["iter"]
        return d["infos"]

    def get_inference_policy(self, device=None):
        self.alg.actor_critic.eval()
        return self.alg.actor_critic.act_inference


class BenchPolicy:
    """What bench.py times next to env.step: PPO.act (actor + critic forward, sampling, log-prob, rollout-storage
    row) writing the next actions in place.  Captured inside the step's CUDA graph, so the storage row and the
    sampling step are the captured ones on every replay (same work, repeated noise)."""

    def __init__(self, env, train_cfg, precision="bf16"):
        from .config import class_to_dict
        cfg = class_to_dict(train_cfg)
        nc = env.num_privileged_obs if env.num_privileged_obs is not None else env.num_obs
        self.ac = ActorCritic(env.num_obs, nc, env.num_actions, device=str(env.device), **cfg["policy"])
        self.alg = PPO(self.ac, device=str(env.device), seed=cfg.get("seed", 1), precision=precision,
                       env_id_offset=getattr(env, "env_id_offset", 0), **cfg["algorithm"])
        self.alg.init_storage(env.num_envs, cfg["runner"]["num_steps_per_env"], [env.num_obs], [env.num_privileged_obs],
                              [env.num_actions])
        self.precision = precision
        path = self.alg.inference_path
        self.kernel_name = {"fp32": "pf_sgemm_kernel chain + pf_act_sample_kernel", "tcgen05": "pf_mlp_tc_kernel",
                            "warp_mma": "pf_mlp_wm_kernel"}[path]
        self.dtype_name = {"fp32": "f32", "tcgen05": "f16 (tcgen05 MLP, fp32 accumulate) + f32 env",
                           "warp_mma": "f16 (register-resident warp-MMA MLP, fp32 accumulate) + f32 env"}[path]
        n_layers = len(self.ac.actor_dims) - 1 + len(self.ac.critic_dims) - 1
        self.launches_per_act = 1 if precision == "bf16" else n_layers + 3      # bf16: ONE fused tcgen05 launch
        self.t = 0
        self.actions = self.alg._actions          # pf_rl_act writes the sampled actions here: step them directly
        # the act kernel is launched right behind the env step's kernels (one CUDA graph): PF_RL_REHEARSE=1 lets it run a dry
        # item under their tail (pf_rl_tc_rehearse).  Off by default since the single-phase program got faster than the
        # rehearsal saves (rough 4,096 envs: 70.3 us per step without, 72.3 with; it was -5 us in round 1).
        # the act launch is captured into the step's graph BEHIND the env step's kernels: it may stage its packed weights before
        # its dependency wait (pf_rl_early_weights; the context cannot see what surrounds a captured launch, so it has to be told)
        self.alg._lib.pf_rl_early_weights(self.alg._ctx, 1)
        if precision == "bf16" and os.environ.get("PF_RL_REHEARSE", "0") != "0":
            self.alg._lib.pf_rl_tc_rehearse(self.alg._ctx, 1)

    def describe(self):
        return (f"ActorCritic actor {self.ac.actor_dims} critic {self.ac.critic_dims} ELU, {self.precision}; "
                "act + evaluate + sample + log-prob + rollout-storage row")

    def act_into(self, env, actions):
        alg = self.alg
        alg.storage.step = self.t
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        a = alg.act(obs, priv if priv is not None else obs)
        if actions.data_ptr() != a.data_ptr():
            actions.copy_(a)
        self.t = (self.t + 1) % alg.storage.num_transitions_per_env


def make_bench_policy(env, train_cfg, precision=None):
    return BenchPolicy(env, train_cfg, precision or os.environ.get("PF_RL_PRECISION", "bf16"))
