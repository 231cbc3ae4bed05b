"""GPU parity of the rsl_rl side (ActorCritic act/evaluate, storage writes, GAE, minibatch shuffle, PPO.update)
through the C-ABI, against oracle/rsl_rl_ref.py.  Tolerances are stated in parity_rl.py."""
import pytest
import torch

import parity_rl
from pointfoot_b200 import rl

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


@pytest.mark.parametrize("hidden,nc,n", [((128, 64, 32), 27, 1), ((128, 64, 32), 27, 15), ((128, 64, 32), 27, 16), ((128, 64, 32), 27, 17),
                                         ((128, 64, 32), 27, 300), ((128, 64, 32), 27, 4097), ((128, 64, 32), 27, 50001),
                                         ((64, 32), 27, 333), ((64, 32), 96, 1000), ((16,), 27, 77), ((32, 16, 16, 16, 16), 100, 130)])
def test_warp_mma_small_net_act_matches_oracle_bf16(hidden, nc, n):
    """Small nets (the flat policy) run register-resident on warp-level MMAs (pf_mlp_wm.cu): same bar as the tcgen05 path,
    row counts around the 16-row task size, input widths that need padding, 1 to 5 hidden layers"""
    ppo, _ = parity_rl.make_pair(8, 27, nc, list(hidden), T=2, precision="bf16")
    try:
        assert ppo.inference_path == "warp_mma"
    finally:
        ppo.close()
    parity_rl.run_act_and_gae(num_envs=n, num_critic_obs=nc, hidden=hidden, T=3, precision="bf16")


@pytest.mark.parametrize("n", [300, 50001])
def test_warp_mma_flat_policy_through_the_layer_switch(n, monkeypatch):
    """The flat policy's four layers are called directly (WmNetArgs.flat3); PF_WM_NO_FLAT3 sends the same nets through the
    per-layer switch every other eligible shape uses: same parity bar, and the same bits as the direct calls"""
    outs = []
    for env in (None, "1"):
        if env is None:
            monkeypatch.delenv("PF_WM_NO_FLAT3", raising=False)
        else:
            monkeypatch.setenv("PF_WM_NO_FLAT3", env)
            parity_rl.run_act_and_gae(num_envs=n, num_critic_obs=27, hidden=(128, 64, 32), T=2, precision="bf16")
        ppo, _ = parity_rl.make_pair(n, 27, 27, [128, 64, 32], T=2, precision="bf16")
        try:
            assert ppo.inference_path == "warp_mma"
            g = torch.Generator(device="cuda").manual_seed(5)
            x = torch.randn(n, 27, device="cuda", generator=g)
            outs.append(ppo.actor_critic.act_inference(x).clone())
        finally:
            ppo.close()
    assert torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("hidden,nc,n", [((128, 64, 32), 27, 300), ((128, 64, 32), 27, 50001)])
def test_tcgen05_path_still_serves_small_nets_when_asked(hidden, nc, n, monkeypatch):
    """PF_RL_WM=0: the compact tcgen05 variant (two CTAs per SM) on the flat nets"""
    monkeypatch.setenv("PF_RL_WM", "0")
    ppo, _ = parity_rl.make_pair(8, 27, nc, list(hidden), T=2, precision="bf16")
    try:
        assert ppo.inference_path == "tcgen05"
    finally:
        ppo.close()
    parity_rl.run_act_and_gae(num_envs=n, num_critic_obs=nc, hidden=hidden, T=3, precision="bf16")


@pytest.mark.parametrize("n,hardwired", [(300, True), (40000, True), (40000, False), (40001, True)])
def test_tcgen05_two_phase_program_matches_oracle(n, hardwired, monkeypatch):
    """The software-pipelined schedule of the same kernel (second half of item m interleaved with the first half of item
    m + 1; pf_mlp_tc.cu build_program_pipe), forced on at every size: prologue + drain only at 300 rows, four to five
    items per CTA at 40,000 (ragged last tile at 40,001) -- with the hard-wired MMA / loader roles of the
    [*, 512, 256, 128, *] nets (tc_mma_fast3, the default for large batches) and on the program interpreter"""
    monkeypatch.setenv("PF_TC_PIPE", "1")
    if not hardwired:
        monkeypatch.setenv("PF_TC_NO_FAST3", "1")
    parity_rl.run_act_and_gae(num_envs=n, num_critic_obs=148, hidden=(512, 256, 128), T=3, precision="bf16")


def test_tcgen05_two_phase_other_shapes_on_the_interpreter(monkeypatch):
    """A three-hidden-layer net that is NOT [*, 512, 256, 128, *]: two layer-0 chunks, one layer-1 chunk -- two-phase program
    on the interpreter (forced; by default such nets take the single-phase program)"""
    monkeypatch.setenv("PF_TC_PIPE", "1")
    parity_rl.run_act_and_gae(num_envs=30000, num_critic_obs=48, hidden=(256, 128, 64), T=2, precision="bf16")


def test_tcgen05_odd_layer_count_many_items_per_cta():
    """Two hidden layers: the output layer accumulates in the TMEM columns the NEXT item's first chunk uses, so the next
    item has to wait for the final epilogue (BAR_ODONE) -- several items per CTA, ragged last tile"""
    parity_rl.run_act_and_gae(num_envs=40001, num_critic_obs=48, hidden=(256, 128), T=2, precision="bf16")


@pytest.mark.parametrize("hidden,nc", [((128, 64, 32), 27), ((512, 256, 128), 148)])
def test_inference_right_behind_a_repack_sees_the_new_weights(hidden, nc):
    """The inference kernels stage their PACKED weights before their programmatic-dependency wait unless the packing kernel may
    be right in front of them (pf_rl_weights_early).  Here it is: load_state_dict -> pack kernel -> act_inference, back to
    back, six times with very different weights; then a second inference call (the early path) on the same weights.
    (The packing kernels never trigger their dependents early, so a dependent only starts once they have exited; the wait
    the bookkeeping adds is what makes their stores formally visible.  A forced always-early build passed this test too.)"""
    torch.manual_seed(5)
    ac = rl.ActorCritic(27, nc, 6, list(hidden), list(hidden), device="cuda:0")
    ppo = rl.PPO(ac, device="cuda:0", precision="bf16")
    ppo.init_storage(3000, 2, [27], [nc], [6])               # creates the context (and the parameter views)
    try:
        obs = torch.randn(3000, 27, device="cuda:0")
        for it in range(6):
            sd = {k: (torch.randn_like(v) * (0.02 + 0.05 * it)) for k, v in ac.state_dict().items()}
            ac.load_state_dict(sd)                       # params_changed(): the pack kernel is now the last thing on the stream
            got = [ac.act_inference(obs), ac.act_inference(obs)]
            x = obs
            n = len(hidden)
            for l in range(n + 1):
                x = x @ sd[f"actor.{2 * l}.weight"].t() + sd[f"actor.{2 * l}.bias"]
                if l < n:
                    x = torch.nn.functional.elu(x)
            # (tolerance relative to the output scale: the point is stale weights -- errors of the size of the outputs --, and
            # the larger weight scales make the outputs large sums with cancellation)
            for g in got:
                torch.testing.assert_close(g, x, rtol=2e-2, atol=2e-2 * float(x.abs().max()), msg=lambda m, it=it: f"iteration {it}: {m}")
    finally:
        ppo.close()


def test_path_selection():
    for hidden, nc, want in (((512, 256, 128), 148, "tcgen05"), ((128, 64, 32), 148, "tcgen05"), ((100, 50), 27, "tcgen05"),
                             ((128, 64, 32), 27, "warp_mma")):
        ppo, _ = parity_rl.make_pair(8, 27, nc, list(hidden), T=2, precision="bf16")
        try:
            assert ppo.inference_path == want, (hidden, nc)
        finally:
            ppo.close()
    ppo, _ = parity_rl.make_pair(8, 27, 27, [128, 64, 32], T=2, precision="fp32")
    assert ppo.inference_path == "fp32"
    ppo.close()


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
    first-minibatch gradient of EVERY parameter tensor within 2e-2 relative L2 of torch autograd, cosine >= 0.999;
    parameters after the whole update within 8 % of the update's own L2 movement (bars and measured values in parity_rl.py)."""
    parity_rl.run_update(num_envs=n, num_critic_obs=nc, hidden=hidden, T=24, precision="bf16")


@pytest.mark.parametrize("hidden,nc,n,t", [((256, 128), 48, 129, 8), ((64, 64, 64, 64), 235, 67, 12), ((200, 48), 27, 300, 4),
                                            ((512, 256, 128), 148, 4096, 24)])
def test_tensor_core_minibatch_gradient_shapes(hidden, nc, n, t):
    """the tcgen05 training GEMMs on other layer shapes (widths that are not multiples of 128, four hidden layers, a ragged
    last row tile, the full 24 x 4,096 minibatch): first-minibatch gradient per parameter tensor against torch autograd"""
    parity_rl.run_update(num_envs=n, num_critic_obs=nc, hidden=hidden, T=t, precision="bf16", first_minibatch_only=True)


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


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_inference_accepts_any_batch_size(precision):
    """ActorCritic.act_inference / evaluate on 1 row, on a ragged count and on more rows than the workspace holds
    (upstream: a plain nn.Sequential call); wrong shapes / dtypes / devices raise instead of reading out of bounds."""
    import torch
    from oracle import rsl_rl_ref
    from pointfoot_b200 import rl
    ppo, oppo = parity_rl.make_pair(64, 27, 148, [512, 256, 128], T=4, precision=precision)
    tol = parity_rl.FP32 if precision == "fp32" else parity_rl.BF16
    try:
        ac, oac = ppo.actor_critic, oppo.actor_critic
        g = torch.Generator().manual_seed(3)
        for rows in (1, 63, 64, 64 * 4 + 5, 1000):          # workspace rows = max(64, 64 * 4 / 4) = 64
            obs, cobs = torch.randn(rows, 27, generator=g), torch.randn(rows, 148, generator=g)
            with torch.no_grad():
                want_a, want_v = oac.act_inference(obs), oac.evaluate(cobs)
            torch.testing.assert_close(ac.act_inference(obs.cuda()).cpu(), want_a, **tol)
            torch.testing.assert_close(ac.evaluate(cobs.cuda()).cpu(), want_v, **tol)
        with pytest.raises(RuntimeError):
            ac.act_inference(torch.zeros(4, 26, device="cuda"))
        with pytest.raises(RuntimeError):
            ac.act_inference(torch.zeros(4, 27))
        with pytest.raises(RuntimeError):
            ac.evaluate(torch.zeros(4, 148, device="cuda", dtype=torch.float64))
        with pytest.raises(RuntimeError):
            ppo.act(torch.zeros(3, 27, device="cuda"), torch.zeros(3, 148, device="cuda"))
        strided = torch.randn(5, 54, generator=g).cuda()[:, ::2]
        with torch.no_grad():
            want = oac.act_inference(strided.cpu().contiguous())
        torch.testing.assert_close(ac.act_inference(strided).cpu(), want, **tol)
    finally:
        ppo.close()


def test_two_shards_sample_the_actions_of_one_big_run():
    """PointFoot stores env_id_offset and OnPolicyRunner hands it to the sampling kernel: two half shards stepped through
    the runner's PPO.act draw exactly actions[off:off+n] of the unsharded run (same parameters, same observations)."""
    import torch
    import common
    from pointfoot_b200 import config, rl, synth
    from pointfoot_b200.env import PointFoot
    N, task = 256, "pointfoot_flat"
    tcfg = config.class_to_dict(config.PointFootFlatCfgPPO())
    tcfg["runner"]["num_steps_per_env"] = 4

    def make(n, off):
        cfg = common.make_cfg(task, n)
        env = PointFoot(cfg, None, None, "cuda:0", True, seed=1, env_id_offset=off)
        assert env.env_id_offset == off and env.seed == 1
        return env, rl.OnPolicyRunner(env, tcfg, None, "cuda:0")
    env, full = make(N, 0)
    obs = torch.randn(N, 27, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    a_full = full.alg.act(obs, obs).clone()
    for off in (0, N // 2):
        e2, half = make(N // 2, off)
        assert half.alg.env_id_offset == off
        half.alg.actor_critic.flat.copy_(full.alg.actor_critic.flat)
        half.alg.actor_critic.params_changed()
        a = half.alg.act(obs[off:off + N // 2].contiguous(), obs[off:off + N // 2].contiguous())
        assert torch.equal(a, a_full[off:off + N // 2])
        half.alg.close(); e2.close()
    full.alg.close(); env.close()
