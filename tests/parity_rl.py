"""Parity driver for the rsl_rl side: the CUDA path (pointfoot_b200.rl over the C-ABI) against the plain-torch
restatement of rsl_rl v1.0.2 (oracle/rsl_rl_ref.py) on the same seeded inputs.  TEST INFRASTRUCTURE.

Tolerances (BASELINE.json north_star: "within a stated fp32 relative tolerance, or a stated looser tolerance
where the MLP runs in TF32/BF16"):
  fp32 path   MLP outputs, log-probs, gradients, updated parameters: rtol 1e-4 / atol 1e-5 (summation order differs
              from ATen's GEMM); GAE returns bit-exact, normalised advantages rtol 1e-5; shuffle indices bit-exact.
  bf16 path   MLP outputs: rtol 2e-2 / atol 2e-2 (bf16 operands, fp32 accumulation).
"""
import numpy as np
import torch

from oracle import rsl_rl_ref
from pointfoot_b200 import rl

FP32 = dict(rtol=1e-4, atol=1e-5)
BF16 = dict(rtol=2e-2, atol=2e-2)
# bf16 tensor-core PPO.update (PF_RL_BF16_TC): operands rounded to bf16 (2^-9 relative), fp32 accumulation.  Per parameter
# tensor the first-minibatch gradient must be within 2e-2 relative L2 of torch autograd with cosine >= 0.999 (measured on
# the B200: worst tensor 5.5e-3 / 0.99999).  Parameters after the 20 Adam steps: within 8 % of the update's own L2 movement
# (measured 6.2 %; the fp32 path sits at 0.2 %).  Adam's first steps move every parameter by ~lr * sign(g), so the 0.5 %
# gradient noise flips the direction of the parameters whose gradient is itself noise-sized; that, not a wrong GEMM, is what
# the remaining 6 % is made of -- the per-tensor gradient bar above is the one that catches a wrong dgrad / wgrad.
GRAD_REL_L2_BF16, GRAD_COS_BF16, PARAM_ERR_BF16 = 2e-2, 0.999, 0.08
ALG = dict(value_loss_coef=1.0, use_clipped_value_loss=True, clip_param=0.2, entropy_coef=0.01, num_learning_epochs=5,
           num_mini_batches=4, learning_rate=1.e-3, schedule="adaptive", gamma=0.99, lam=0.95, desired_kl=0.01,
           max_grad_norm=1.)


def make_pair(num_envs, num_obs, num_critic_obs, hidden, T=24, seed=1, precision="fp32", alg=None, num_actions=6):
    """(cuda PPO, oracle PPO) with identical parameters."""
    alg = dict(ALG, **(alg or {}))
    ac = rl.ActorCritic(num_obs, num_critic_obs, num_actions, hidden, hidden, device="cuda:0")
    ppo = rl.PPO(ac, device="cuda:0", seed=seed, precision=precision, **alg)
    ppo.init_storage(num_envs, T, [num_obs], [num_critic_obs], [num_actions])
    g = torch.Generator().manual_seed(seed)
    ac.reset_parameters(generator=g)
    with torch.no_grad():                       # a non-trivial std so that its gradient is exercised
        ac.std.copy_(torch.linspace(0.6, 1.1, num_actions))
    oac = rsl_rl_ref.ActorCritic(num_obs, num_critic_obs, num_actions, hidden, hidden)
    oac.load_state_dict({k: v.detach().cpu().clone() for k, v in ac.state_dict().items()})
    oac.sample_seed = seed
    oppo = rsl_rl_ref.PPO(oac, **alg)
    oppo.seed = seed
    oppo.init_storage(num_envs, T, [num_obs], [num_critic_obs], [num_actions])
    return ppo, oppo


def flat_params(oac):
    return torch.cat([p.detach().flatten() for p in oac.parameters()])


def rollout(ppo, oppo, T, seed=0, obs_scale=1.0):
    """Fill both storages through act / process_env_step with the same synthetic env outputs."""
    g = torch.Generator().manual_seed(1000 + seed)
    N = ppo.storage.num_envs
    no, nc = ppo.actor_critic.num_actor_obs, ppo.actor_critic.num_critic_obs
    out = []
    for t in range(T):
        obs = torch.randn(N, no, generator=g) * obs_scale
        cobs = torch.cat([obs, torch.randn(N, nc - no, generator=g)], 1) if nc > no else obs
        rew = torch.randn(N, generator=g) * 0.1
        dones = torch.rand(N, generator=g) < 0.05
        touts = dones & (torch.rand(N, generator=g) < 0.5)
        a = ppo.act(obs.cuda(), cobs.cuda()).clone()
        with torch.no_grad():
            oa = oppo.act(obs, cobs)
        out.append((a.cpu(), oa))
        ppo.process_env_step(rew.cuda(), dones.cuda(), {"time_outs": touts.cuda()})
        oppo.process_env_step(rew, dones, {"time_outs": touts})
    last = torch.randn(N, nc, generator=g)
    return out, last


def check_storage(ppo, oppo, tol, names=("observations", "privileged_observations", "actions", "values", "actions_log_prob",
                                         "mu", "sigma", "rewards", "dones")):
    for n in names:
        a, b = getattr(ppo.storage, n), getattr(oppo.storage, n)
        if a is None or b is None:
            assert a is None and b is None, n
            continue
        if n in ("observations", "privileged_observations", "dones", "sigma"):
            assert torch.equal(a.cpu(), b), f"storage.{n} not bit-exact"
        else:
            torch.testing.assert_close(a.cpu(), b, msg=lambda m, n=n: f"storage.{n}: {m}", **tol)


def run_act_and_gae(num_envs=256, num_obs=27, num_critic_obs=148, hidden=(512, 256, 128), T=24, precision="fp32", seed=1):
    tol = FP32 if precision == "fp32" else BF16
    ppo, oppo = make_pair(num_envs, num_obs, num_critic_obs, list(hidden), T, seed, precision)
    try:
        acts, last = rollout(ppo, oppo, T, seed)
        for t, (a, oa) in enumerate(acts):
            torch.testing.assert_close(a, oa, msg=lambda m, t=t: f"actions step {t}: {m}", **tol)
        check_storage(ppo, oppo, tol)
        if precision == "fp32":
            # make the GAE inputs identical so that the reverse scan itself can be checked bit for bit
            for n in ("rewards", "values"):
                getattr(ppo.storage, n).copy_(getattr(oppo.storage, n).cuda())
        ppo.compute_returns(last.cuda())
        with torch.no_grad():
            oppo.compute_returns(last)
        if precision == "fp32":
            # last_values come from the two critics (tolerance); everything before the last step's bootstrap is exact
            torch.testing.assert_close(ppo.storage.returns.cpu(), oppo.storage.returns, **tol)
            torch.testing.assert_close(ppo.storage.advantages.cpu(), oppo.storage.advantages, rtol=1e-4, atol=1e-4)
        else:
            torch.testing.assert_close(ppo.storage.returns.cpu(), oppo.storage.returns, **tol)
    finally:
        ppo.close()
    return True


def run_gae_exact(num_envs=1000, T=24, seed=3):
    """compute_returns alone on identical inputs: returns bit-exact, normalised advantages to 1e-5."""
    ppo, oppo = make_pair(num_envs, 27, 27, [32, 16, 8], T, seed)
    try:
        g = torch.Generator().manual_seed(seed)
        for n, mk in (("rewards", lambda: torch.randn(T, num_envs, 1, generator=g)),
                      ("values", lambda: torch.randn(T, num_envs, 1, generator=g)),
                      ("dones", lambda: (torch.rand(T, num_envs, 1, generator=g) < 0.1).to(torch.uint8))):
            v = mk()
            getattr(oppo.storage, n).copy_(v)
            getattr(ppo.storage, n).copy_(v.cuda())
        last = torch.randn(num_envs, 1, generator=g)
        oppo.storage.compute_returns(last, oppo.gamma, oppo.lam)
        lib, ctx = ppo._lib, ppo._ctx
        import ctypes as C
        from pointfoot_b200 import capi
        lv = last.flatten().cuda().contiguous()
        st = rl._stream(ppo.device)
        capi.check(lib.pf_rl_gae(ctx, C.byref(ppo.storage.c), rl._ptr(lv), rl._ptr(ppo._gae_stats), st))
        raw = ppo.storage.advantages.cpu().clone()
        capi.check(lib.pf_rl_normalize_advantages(ctx, C.byref(ppo.storage.c), rl._ptr(ppo._gae_stats), st))
        torch.cuda.synchronize()
        assert torch.equal(ppo.storage.returns.cpu(), oppo.storage.returns), "GAE returns not bit-exact"
        assert torch.equal(raw, oppo.storage.returns - oppo.storage.values), "raw advantages not bit-exact"
        torch.testing.assert_close(ppo.storage.advantages.cpu(), oppo.storage.advantages, rtol=1e-5, atol=1e-6)
    finally:
        ppo.close()
    return True


def run_permutation(n_envs=4096, T=24, iteration=5, seed=1, num_mini_batches=4):
    ppo, _ = make_pair(n_envs, 27, 27, [32, 16, 8], T, seed, alg=dict(num_mini_batches=num_mini_batches))
    try:
        import ctypes as C
        from pointfoot_b200 import capi
        capi.check(ppo._lib.pf_rl_make_permutation(ppo._ctx, C.c_int64(iteration), rl._stream(ppo.device)))
        torch.cuda.synchronize()
        n = (n_envs * T // num_mini_batches) * num_mini_batches
        want = rsl_rl_ref.feistel_permutation(n, seed, iteration)
        # the ctx keeps the permutation; read it back through a minibatch-sized gather of arange
        import numpy as np
        got = torch.empty(n, dtype=torch.int32, device="cuda")
        capi.check(ppo._lib.pf_rl_get_permutation(ppo._ctx, rl._ptr(got), rl._stream(ppo.device)))
        torch.cuda.synchronize()
        assert np.array_equal(got.cpu().numpy().astype(np.int64), want), "minibatch permutation differs from the contract"
        assert len(set(want.tolist())) == n
    finally:
        ppo.close()
    return True


def run_update(num_envs=512, num_obs=27, num_critic_obs=148, hidden=(512, 256, 128), T=24, seed=1, precision="fp32",
               first_minibatch_only=False):
    """PPO.update on identical storages: gradient of the first minibatch, then the parameters after all
    num_learning_epochs x num_mini_batches steps, the adaptive learning rate and the mean losses."""
    tol = FP32 if precision == "fp32" else BF16
    ppo, oppo = make_pair(num_envs, num_obs, num_critic_obs, list(hidden), T, seed, precision)
    try:
        _, last = rollout(ppo, oppo, T, seed)
        with torch.no_grad():
            oppo.compute_returns(last)
        # identical storages: take the oracle's
        for n in ("actions", "values", "actions_log_prob", "mu", "sigma", "rewards", "returns", "advantages"):
            getattr(ppo.storage, n).copy_(getattr(oppo.storage, n).cuda())
        import ctypes as C
        from pointfoot_b200 import capi
        # ---- first minibatch: gradient ----
        capi.check(ppo._lib.pf_rl_make_permutation(ppo._ctx, C.c_int64(ppo.iteration), rl._stream(ppo.device)))
        ppo.minibatch_grad(0)
        torch.cuda.synchronize()
        g_cuda = ppo.grads.cpu().clone()
        stats = ppo.loss_stats.cpu().clone()
        p0 = flat_params(oppo.actor_critic).clone()
        # the oracle's first minibatch: a full update on a deep copy, reading the hooks at its first optimizer step
        import copy
        oc = copy.deepcopy(oppo)
        first = {}
        orig_step = oc.optimizer.step

        def grab(*a, **k):
            if "g" not in first:
                first["g"] = oc.last_grad.clone()
                first["kl"] = oc.last_kl[0]
            return orig_step(*a, **k)
        oc.optimizer.step = grab
        oc.update()
        scale = first["g"].abs().max().item()
        if precision == "fp32":
            torch.testing.assert_close(g_cuda, first["g"], rtol=tol["rtol"] * 10, atol=max(tol["atol"], tol["rtol"] * scale),
                                       msg=lambda m: f"first-minibatch gradient: {m}")
        # per parameter tensor (std, every weight, every bias): relative L2 error and cosine against autograd -- a wrong
        # dgrad / wgrad of ONE small layer cannot hide behind the large ones
        rel_bar, cos_bar = (1e-3, 0.999999) if precision == "fp32" else (GRAD_REL_L2_BF16, GRAD_COS_BF16)
        lay, _ = ppo.actor_critic.layout()
        worst = (0.0, 1.0, "")
        for key, off, shape in lay:
            n = int(torch.Size(shape).numel())
            a, b = g_cuda[off:off + n].double(), first["g"][off:off + n].double()
            nb = b.norm().item()
            if nb < 1e-12:
                assert a.norm().item() < 1e-9, key
                continue
            rel = (a - b).norm().item() / nb
            cos = (a @ b).item() / (a.norm().item() * nb + 1e-300)
            if rel > worst[0]:
                worst = (rel, cos, key)
            assert rel <= rel_bar and cos >= cos_bar, f"gradient of {key}: rel-L2 {rel:.3e} (bar {rel_bar}), cosine {cos:.6f} (bar {cos_bar})"
        print(f"[{precision}] first-minibatch gradient, worst tensor {worst[2]}: rel-L2 {worst[0]:.3e} cosine {worst[1]:.6f}")
        kl_cuda = (stats[3] / stats[4]).item()
        assert abs(kl_cuda - first["kl"]) <= 1e-4 + 1e-3 * abs(first["kl"]), (kl_cuda, first["kl"])
        if first_minibatch_only:
            return True
        # ---- the whole update ----
        vl, sl = ppo.update()
        ovl, osl = oppo.update()
        torch.cuda.synchronize()
        p_cuda, p_ref = ppo.actor_critic.flat.cpu(), flat_params(oppo.actor_critic)
        # Adam's normalised step turns rounding noise on near-zero gradients into +-lr moves, so the bar is on the
        # update as a whole (relative L2), not per element
        moved = (p_ref - p0).norm().item()
        err = (p_cuda - p_ref).norm().item()
        print(f"[{precision}] parameters after the update: |err| {err:.3e} = {err / moved:.4f} of the movement {moved:.3e}")
        assert err <= (0.02 if precision == "fp32" else PARAM_ERR_BF16) * moved, f"parameters after update: |err| {err:.3e} vs |movement| {moved:.3e}"
        assert abs(ppo.learning_rate - oppo.learning_rate) <= 1e-12 + 1e-9 * oppo.learning_rate, (ppo.learning_rate, oppo.learning_rate)
        lr_, la_ = (1e-3, 1e-5) if precision == "fp32" else (2e-2, 2e-4)      # mean losses: bf16 operands move them by ~1 %
        assert abs(vl - ovl) <= lr_ * abs(ovl) + la_ and abs(sl - osl) <= lr_ * abs(osl) + la_, (vl, ovl, sl, osl)
    finally:
        ppo.close()
    return True


def smoke_check():
    run_act_and_gae(num_envs=64, T=4)
    run_gae_exact(num_envs=64, T=8)
    print("rl parity: act / storage / GAE ok")
    # the tensor-core kernels (tcgen05): the fused act kernel and one minibatch of PPO.update on the bf16 GEMMs
    run_act_and_gae(num_envs=256, num_critic_obs=148, hidden=(512, 256, 128), T=2, precision="bf16")
    run_update(num_envs=128, num_critic_obs=148, hidden=(512, 256, 128), T=8, precision="bf16", first_minibatch_only=True)
    print("rl parity: tcgen05 act + tcgen05 minibatch gradient ok")
