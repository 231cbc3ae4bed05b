"""CPU checks of the rsl_rl oracle restatement (oracle/rsl_rl_ref.py) and of the host-side contracts."""
import numpy as np
import pytest
import torch

from oracle import ref_harness, rsl_rl_ref


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


def test_tensor_core_mlp_schedule_is_built_and_dry_run_on_the_host():
    """pf_rl_tc_plan (no device): shapes the fused tcgen05 MLP kernel accepts, and the structure of its static schedule
    for the PointFoot critic [148, 512, 256, 128, 1]: 4 layer-0 chunks + 4 + 2 + 1 K-slices = 11 MMA groups, 8 epilogues,
    every chunk buffer hand-over guarded by a barrier."""
    import ctypes as C
    import os
    import re
    from pointfoot_b200 import capi
    lib = capi.load()

    def plan(dims, cap=16384):
        arr = (C.c_int32 * len(dims))(*dims)
        buf = C.create_string_buffer(cap)
        return lib.pf_rl_tc_plan(arr, len(dims) - 1, buf, cap), buf.value.decode()

    for ok in ([148, 512, 256, 128, 1], [27, 128, 64, 32, 6], [48, 256, 128, 6], [235, 64, 64, 64, 12], [27, 6], [27, 1024, 256, 6]):
        assert plan(ok)[0] == 0, ok
    for bad in ([300, 64, 6], [27, 64, 512, 6], [27, 1100, 32, 1]):      # input > 256, later layer > 256, first layer > 1024
        assert plan(bad)[0] == 6, bad
    rc, text = plan([148, 512, 256, 128, 1])
    groups = re.findall(r"^G\d+ ph=\d+ layer=(\d+) a_src=(\d+) nt=(\d+) kc=(\d+) kt=(\d+) col=(\d+) bias=(\d+) kt0=\d+ wait=(-?\d+)@", text, re.M)
    epis = re.findall(r"^E\d+ ph=\d+ col=(\d+) nt=(\d+) n0=(\d+) acc=(\d+)@(\d+) dst=(-?\d+)", text, re.M)
    assert len(groups) == 11 and len(epis) == 8
    assert [g[0] for g in groups].count("0") == 4 and [g[0] for g in groups].count("1") == 4
    assert all(int(g[2]) * int(g[4]) * 2 <= 32768 for g in groups)            # a weight tile fits a ring slot
    assert all(int(g[7]) >= 0 for g in groups if g[1] != "0")                 # chunk consumers wait for their producer
    assert [e[5] for e in epis][-1] == "-1" and all(e[5] in "01" for e in epis[:-1])
    blob = int(re.search(r"blob_bytes=(\d+)", text).group(1))
    assert blob == 2 * ((160 + 16) * 512 + (512 + 16) * 256 + (256 + 16) * 128 + (128 + 16) * 16)
    # the TWO-PHASE program of the same net (large batches): the same 11 groups and 8 epilogues, re-ordered so that the second
    # half of an item (phase 1: layer-1 epilogues, layers 2 and 3) interleaves with the first half of the next one; the
    # same packed blob; replayed on the host over three items by check_program (rc 6 if it cannot finish)
    os.environ["PF_TC_PLAN_PIPE"] = "1"
    try:
        rc2, text2 = plan([148, 512, 256, 128, 1])
        assert rc2 == 0
        g2 = re.findall(r"^G\d+ ph=(\d+) layer=(\d+) a_src=(\d+) nt=(\d+) kc=(\d+) kt=(\d+) col=(\d+)", text2, re.M)
        e2 = re.findall(r"^E\d+ ph=(\d+) col=(\d+) nt=(\d+) n0=(\d+) acc=(\d+)@(\d+) dst=(-?\d+)", text2, re.M)
        assert len(g2) == 11 and len(e2) == 8
        assert [e[0] for e in e2] == ["1", "0"] * 4                               # epilogue steps alternate between the two items
        assert sorted((g[1], g[3], g[4]) for g in g2) == sorted((g[0], g[2], g[3]) for g in groups)
        assert int(re.search(r"blob_bytes=(\d+)", text2).group(1)) == blob
        cols = {(g[0], g[1]): int(g[6]) for g in g2}
        assert cols[("0", "0")] == 0 and cols[("0", "1")] == 256 and cols[("1", "2")] == 128 and cols[("1", "3")] == 128
        for dims in ([27, 512, 256, 128, 6], [48, 256, 128, 64, 6], [48, 1024, 256, 128, 12]):
            assert plan(dims)[0] == 0, dims
        for dims in ([27, 128, 64, 32, 6], [48, 256, 128, 6], [27, 512, 256, 256, 6]):     # one layer-0 chunk / not 3 hidden layers / layer 3 too wide
            assert plan(dims)[0] == 6, dims
    finally:
        del os.environ["PF_TC_PLAN_PIPE"]


def test_export_policy_as_jit_from_a_reference_style_checkpoint(tmp_path):
    """a state dict with the upstream keys (here: the oracle's ActorCritic) exports to the TorchScript
    actor the reference writes (helpers.py:180-190), and the file reproduces act_inference."""
    import torch
    from oracle import rsl_rl_ref
    from pointfoot_b200 import helpers
    torch.manual_seed(3)
    ac = rsl_rl_ref.ActorCritic(27, 148, 6, [64, 32, 16], [64, 32, 16])

    class Shim:                      # what export needs from either stack's ActorCritic
        is_recurrent = False
        def state_dict(self):
            return ac.state_dict()
    file = helpers.export_policy_as_jit(Shim(), str(tmp_path))
    m = torch.jit.load(file)
    x = torch.randn(5, 27)
    assert torch.equal(m(x), ac.act_inference(x))


def test_logger_records_like_the_reference(tmp_path, capsys):
    """legged_gym/utils/logger.py:36-133, data side: same keys, the rewards weighted by the episode count, the headless plot"""
    import numpy as np
    import torch
    from pointfoot_b200.logger import Logger
    lg = Logger(0.02)
    for i in range(5):
        lg.log_states({"dof_pos": 0.1 * i, "base_vel_x": torch.tensor(float(i)), "contact_forces_z": np.array([1.0 * i, 2.0])})
    lg.log_rewards({"rew_tracking_lin_vel": torch.tensor(0.5), "terrain_level": torch.tensor(3.0)}, 2)
    lg.log_rewards({"rew_tracking_lin_vel": torch.tensor(0.25)}, 2)
    assert lg.num_episodes == 4 and list(lg.rew_log) == ["rew_tracking_lin_vel"] and lg.rew_log["rew_tracking_lin_vel"] == [1.0, 0.5]
    lg.print_rewards()
    out = capsys.readouterr().out
    assert " - rew_tracking_lin_vel: 0.375" in out and "Total number of episodes: 4" in out
    path = lg.plot_states(str(tmp_path / "states"))
    z = np.load(str(tmp_path / "states.npz"))
    assert path.startswith(str(tmp_path)) and z["contact_forces_z"].shape == (5, 2) and np.allclose(z["time"], np.linspace(0, 0.1, 5))
    if ref_harness.reference_available():
        # same public methods as the reference class (read from its source: matplotlib is not installed here)
        import ast
        src = open(ref_harness.REFERENCE_ROOT + "/legged_gym/utils/logger.py").read()
        cls = [n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == "Logger"][0]
        names = {f.name for f in cls.body if isinstance(f, ast.FunctionDef) and not f.name.startswith("_")}
        assert names <= set(dir(Logger))
    lg.reset()
    assert not lg.state_log and not lg.rew_log
