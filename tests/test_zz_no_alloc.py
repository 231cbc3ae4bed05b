"""pf_rl_gae_noalloc (include/pointfoot_b200.h): the GAE reverse scan for callers that cannot allocate device memory --
no pf_rl_ctx, the reduction's scratch in a static device array inside the library.  It has no counterpart in the
reference; the bar is the existing entry point pf_rl_gae on the same
inputs: returns, raw advantages and the three statistics BIT-EXACT (the tests' bar for GAE, tests/parity_rl.py), on two
calls in a row that share the array.  Collected last on purpose (pytest -x must not hide other tests behind it)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ENVS = 1048576          # PF_GAE_NOALLOC_MAX_ENVS


def _inputs(n, t, seed):
    g = torch.Generator().manual_seed(seed)
    rewards = torch.randn(t, n, 1, generator=g)
    values = torch.randn(t, n, 1, generator=g)
    dones = (torch.rand(t, n, 1, generator=g) < 0.1).to(torch.uint8)
    last = torch.randn(n, generator=g)
    return rewards.cuda(), dones.cuda(), values.cuda(), last.cuda()


def _existing(n, t, rewards, dones, values, last, gamma, lam):
    """pf_rl_gae through a pf_rl_ctx (pointfoot_b200.rl.PPO owns one) -> returns, raw advantages, stats"""
    from pointfoot_b200 import capi, rl
    ac = rl.ActorCritic(27, 27, 6, [32, 16, 8], [32, 16, 8], device="cuda:0")
    ppo = rl.PPO(ac, device="cuda:0", gamma=gamma, lam=lam, num_mini_batches=1, num_learning_epochs=1)
    ppo.init_storage(n, t, [27], [27], [6])
    try:
        st = ppo.storage
        st.rewards.copy_(rewards); st.dones.copy_(dones); st.values.copy_(values)
        stats = torch.zeros(3, dtype=torch.float64, device="cuda")
        capi.check(ppo._lib.pf_rl_gae(ppo._ctx, C.byref(st.c), C.c_void_p(last.data_ptr()), C.c_void_p(stats.data_ptr()),
                                      rl._stream(ppo.device)), "pf_rl_gae")
        torch.cuda.synchronize()
        return st.returns.clone(), st.advantages.clone(), stats.clone()
    finally:
        ppo.close()


def _noalloc(lib, n, t, rewards, dones, values, last, gamma, lam):
    returns = torch.full((t, n, 1), float("nan"), device="cuda")
    adv = torch.full((t, n, 1), float("nan"), device="cuda")
    stats = torch.full((3,), float("nan"), dtype=torch.float64, device="cuda")
    p = lambda x: C.c_void_p(x.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    before = torch.cuda.memory_allocated()
    rc = lib.pf_rl_gae_noalloc(p(rewards), p(dones), p(values), p(last), C.c_int32(n), C.c_int32(t), C.c_float(gamma),
                               C.c_float(lam), p(returns), p(adv), p(stats), stream)
    assert rc == 0, lib.pf_last_error()
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    return returns, adv, stats


def test_gae_noalloc_matches_pf_rl_gae_on_two_calls_in_a_row():
    from pointfoot_b200 import capi
    lib = capi.load_rl()
    gamma, lam = 0.99, 0.95
    # the second call is SMALLER than the first (fewer CTAs): partials the first call left in the shared array beyond the
    # second call's CTAs must not leak into its statistics; the third repeats the first shape with other data
    for n, t, seed in ((5000, 24, 1), (777, 7, 2), (5000, 24, 3)):
        rewards, dones, values, last = _inputs(n, t, seed)
        got = _noalloc(lib, n, t, rewards, dones, values, last, gamma, lam)
        want = _existing(n, t, rewards, dones, values, last, gamma, lam)
        for name, a, b in zip(("returns", "advantages", "stats"), got, want):
            assert torch.equal(a, b), f"pf_rl_gae_noalloc {name} differs from pf_rl_gae (N={n}, T={t})"
        assert float(got[2][2]) == n * t


def test_gae_noalloc_rejects_sizes_beyond_the_static_array():
    from pointfoot_b200 import capi
    lib = capi.load_rl()
    x = torch.zeros(8, device="cuda")
    d = torch.zeros(8, dtype=torch.uint8, device="cuda")
    s = torch.zeros(3, dtype=torch.float64, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = lib.pf_rl_gae_noalloc(p(x), p(d), p(x), p(x), C.c_int32(MAX_ENVS + 1), C.c_int32(1), C.c_float(0.99), C.c_float(0.95),
                               p(x), p(x), p(s), None)
    assert rc == 8 and b"PF_GAE_NOALLOC_MAX_ENVS" in lib.pf_last_error()
    assert lib.pf_rl_gae_noalloc(None, p(d), p(x), p(x), C.c_int32(8), C.c_int32(1), C.c_float(0.99), C.c_float(0.95),
                                 p(x), p(x), p(s), None) == 1
