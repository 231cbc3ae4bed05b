"""Out-of-bounds writes: every buffer the kernels write sits between 1 KB guard bands that must come back untouched, at
env counts that exercise the tails (1, 33, 4097) as well as full tiles (4096).  The stand-in for a compute-sanitizer
memcheck run, which this GPU pool refuses (profiles/r2_sanitizer_unavailable.txt)."""
import ctypes as C

import pytest
import torch

import common
from pointfoot_b200 import capi, synth

pytestmark = pytest.mark.gpu
SIZES = [1, 33, 4096, 4097]


@pytest.mark.parametrize("n", SIZES)
@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
def test_env_step_kernels_stay_inside_their_buffers(task, n):
    import parity
    state = synth.make_state(n, task, seed=3, adversarial=True)
    env = parity.EnvUnderTest(task, state, "cuda", common.push_every_step_2)
    rz = common.RedZone()
    try:
        for name in capi.ENV_BUFFER_FIELDS:
            if env.t.get(name) is not None and name not in ("height_samples", "terrain_origins", "terrain_types"):
                env.t[name] = rz.wrap(env.t[name], name)
                setattr(env.b, name, env.t[name].data_ptr())
        env.t["push_forces"] = rz.wrap(env.t["push_forces"], "push_forces")
        for i in range(3):
            env.step(synth.make_state(n, task, seed=10 + i)["actions"].cuda())
        capi.check(env.lib.pf_reset_all(env.ctx, C.byref(env.b), C.c_int64(7), None), "pf_reset_all")
        torch.cuda.synchronize()
        assert rz.check() > 25
    finally:
        env.close()


@pytest.mark.parametrize("n", SIZES)
def test_legged_robot_kernels_stay_inside_their_buffers(n):
    from pointfoot_b200 import config
    from pointfoot_b200.legged_robot import LeggedRobot
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = n
    cfg.domain_rand.push_interval_s = 0.04
    env = LeggedRobot(cfg, None, None, "cuda:0", True, state=synth.make_lr_state(n, seed=3, adversarial=True))
    rz = common.RedZone()
    for name in capi.LR_BUFFER_FIELDS:
        t = env._buf_refs[name]
        if t is not None and name not in ("height_samples", "terrain_origins", "terrain_types", "noise_scale_vec", "height_points_xy"):
            env._buf_refs[name] = rz.wrap(t, name)
            setattr(env._bufs, name, env._buf_refs[name].data_ptr())
    env.reset()
    for _ in range(3):
        env.step(torch.randn(n, 12, device="cuda"))
    torch.cuda.synchronize()
    assert rz.check() > 20


@pytest.mark.parametrize("n", SIZES)
@pytest.mark.parametrize("precision,hidden", [("fp32", (64, 32)), ("bf16", (512, 256, 128)), ("bf16", (128, 64, 32))])
def test_rollout_kernels_stay_inside_the_storage(precision, hidden, n):
    """PPO.act (both nets, sample, storage row), process_env_step, GAE, minibatch gathers: the [T, N, .] storage tensors"""
    from pointfoot_b200 import rl
    T = 3
    ac = rl.ActorCritic(27, 148, 6, hidden, hidden, device="cuda:0")
    ppo = rl.PPO(ac, num_learning_epochs=1, num_mini_batches=2, precision=precision, device="cuda:0")
    ppo.init_storage(n, T, [27], [148], [6])
    rz = common.RedZone()
    st = ppo.storage
    for name in capi.RL_STORAGE_FIELDS:
        t = getattr(st, name)
        if t is not None:
            setattr(st, name, rz.wrap(t, name))
            setattr(st.c, name, getattr(st, name).data_ptr())
    try:
        for t in range(T):
            ppo.act(torch.randn(n, 27, device="cuda"), torch.randn(n, 148, device="cuda"))
            ppo.process_env_step(torch.randn(n, device="cuda") * 0.1, torch.rand(n, device="cuda") < 0.1,
                                 {"time_outs": torch.zeros(n, dtype=torch.bool, device="cuda")})
        ppo.compute_returns(torch.randn(n, 148, device="cuda"))
        ppo.update()
        torch.cuda.synchronize()
        assert rz.check() >= 10
    finally:
        ppo.close()


@pytest.mark.parametrize("n", SIZES)
def test_actuator_and_terrain_kernels_stay_inside_their_buffers(n):
    from pointfoot_b200.anymal import ActuatorNetwork
    rz = common.RedZone()
    net = ActuatorNetwork(n, 12, 0.5, torch.zeros(1, 12, device="cuda"), "cuda:0")
    net.weights = (torch.randn(capi.ACTUATOR_NET_WEIGHTS, device="cuda") * 0.3).contiguous()
    for k in ("sea_hidden_state", "sea_cell_state", "torques"):
        setattr(net, k, rz.wrap(getattr(net, k), k))
    dof = torch.randn(n * 12, 2, device="cuda")
    for _ in range(2):
        net.compute_torques(torch.randn(n, 12, device="cuda"), dof)
    torch.cuda.synchronize()
    assert rz.check() == 3
    # terrain: the map, the origins and the mesh (sizes do not depend on n; odd cell counts / a 1.5 m border)
    import parity_terrain
    t = parity_terrain.make_impl("terrain_all_curriculum.npz", "cuda:0")
    rz2 = common.RedZone()
    hf, org = rz2.wrap(t.height_field_raw, "height_field"), rz2.wrap(t.env_origins_device, "env_origins")
    t._call_generate(t._params, t._subs, hf, org)
    R, Cn = hf.shape
    v, tri = rz2.wrap(t.vertices, "vertices"), rz2.wrap(t.triangles, "triangles")
    t._call_trimesh(hf, R, Cn, t.cfg.horizontal_scale, t.cfg.vertical_scale, t.cfg.slope_treshold, v, tri)
    torch.cuda.synchronize()
    assert rz2.check() == 4 and torch.equal(hf, t.height_field_raw) and torch.equal(v, t.vertices)
