"""GPU parity of the env kernels, through the C-ABI: against the committed outputs of the unmodified reference
(tests/golden) and against the oracle on seeded synthetic states, including ragged / edge sizes."""
import pytest
import torch

import parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fname", sorted(parity.GOLDEN))
def test_cuda_matches_reference_golden(fname):
    parity.golden_impl(fname, device="cuda", backend="cuda")


@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
@pytest.mark.parametrize("n", [1, 31, 64, 1000, 4096])
def test_cuda_matches_oracle(task, n):
    parity.run(task, n, steps=3)


def test_cuda_matches_oracle_187_point_grid():
    import common
    parity.run("pointfoot_rough", 2048, steps=2, mutate=common.grid187)


def test_cuda_matches_oracle_lr_command_resampling():
    parity.run("pointfoot_rough", 1024, steps=3, resample_commands=True)


def test_cuda_matches_oracle_large_multi_epb():
    # 20000 envs -> 64 envs per CTA, 70000 -> 128: covers every CTA geometry incl. a ragged last CTA
    parity.run("pointfoot_rough", 20000, steps=2)
    parity.run("pointfoot_flat", 70001, steps=2)


@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
def test_cuda_matches_oracle_at_full_size(task):
    """BASELINE.json's largest per-GPU size (262,144 envs, 32 envs per warp): the same bars as the small cases, on
    every env -- bit-exact flags / indices / reset ids, rtol 1e-5 on observations and rewards."""
    out = parity.run(task, 262144, steps=2, adversarial=False)
    assert out[-1]["num_resets"] > 0                      # the synthetic state makes ~0.3 % of the envs reset per step


def test_envs_per_warp_geometries_agree_bit_for_bit(monkeypatch):
    """The step kernel's result must not depend on how many envs a warp owns (PF_EPW overrides the size heuristic)."""
    from pointfoot_b200 import synth
    state = synth.make_state(3000, "pointfoot_rough", seed=7)
    outs = []
    for epw in ("8", "16", "32"):
        monkeypatch.setenv("PF_EPW", epw)
        e = parity.EnvUnderTest("pointfoot_rough", state, "cuda", seed=7)
        e.step(state["actions"])
        e.step(state["actions"] * 0.5)
        outs.append({k: e.t[k].clone() for k in ("obs_buf", "priv_obs_buf", "rew_buf", "reset_buf", "episode_sums", "reset_ids",
                                                   "terrain_levels", "root_states", "episode_stats")})
        e.close()
    for o in outs[1:]:
        for k, v in o.items():
            if k == "episode_stats":                       # per-warp partial sums: fixed tree per geometry, not across them
                torch.testing.assert_close(v, outs[0][k], rtol=1e-5, atol=1e-6)
            else:
                assert torch.equal(v, outs[0][k]), k


def test_reward_statements_and_cta_sizes_agree_bit_for_bit(monkeypatch):
    """The three statements of the step (pointfoot term set compiled in, table-driven sum for any subset in dir()
    order, order-agnostic loop) give the same bits; so does every CTA size the launch heuristic can pick."""
    from pointfoot_b200 import synth
    state = synth.make_state(3000, "pointfoot_rough", seed=5)
    outs = []
    for env in ({}, {"PF_GENERIC_REWARDS": "1"}, {"PF_TABLE_REWARDS": "1"}, {"PF_BLOCK": "32"}, {"PF_BLOCK": "128", "PF_EPW": "4"}):
        for k in ("PF_GENERIC_REWARDS", "PF_TABLE_REWARDS", "PF_BLOCK", "PF_EPW"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        e = parity.EnvUnderTest("pointfoot_rough", state, "cuda", seed=5)
        e.step(state["actions"])
        e.step(state["actions"] * 0.5)
        outs.append({k: e.t[k].clone() for k in ("obs_buf", "priv_obs_buf", "rew_buf", "reset_buf", "episode_sums", "reset_ids",
                                                   "terrain_levels", "root_states", "last_root_vel", "last_actions")})
        e.close()
    for o in outs[1:]:
        for k, v in o.items():
            assert torch.equal(v, outs[0][k]), k


def test_unaligned_state_tensor_takes_the_generic_kernel():
    """The specialised kernels read pairs of the per-env state as 64-bit words; a caller-owned tensor that is only
    4-byte aligned (a slice) must still work: the launch falls back to the generic kernel, same bits."""
    from pointfoot_b200 import synth
    state = synth.make_state(777, "pointfoot_flat", seed=13)
    a = parity.EnvUnderTest("pointfoot_flat", state, "cuda", seed=13)
    b = parity.EnvUnderTest("pointfoot_flat", state, "cuda", seed=13)
    odd = torch.zeros(b.t["last_actions"].numel() + 1, device="cuda")[1:].view_as(b.t["last_actions"])
    assert odd.data_ptr() % 8 == 4
    odd.copy_(b.t["last_actions"])
    b.t["last_actions"] = odd
    b.b.last_actions = odd.data_ptr()
    for k in range(3):
        a.step(state["actions"] * (1.0 - 0.3 * k)); b.step(state["actions"] * (1.0 - 0.3 * k))
    for name in ("obs_buf", "rew_buf", "reset_buf", "episode_sums", "last_actions", "feet_air_time", "contact_filt", "root_states"):
        assert torch.equal(a.t[name], b.t[name]), name
    a.close(); b.close()


def test_split_tail_equals_plain_step():
    """pf_post_physics_step_begin / _join (reset ids, count, episode statistics on the context's side stream, beside
    whatever the caller queues next) leaves exactly what pf_post_physics_step leaves, over several steps."""
    from pointfoot_b200 import synth
    state = synth.make_state(5000, "pointfoot_rough", seed=9)
    envs = []
    for split in (False, True):
        e = parity.EnvUnderTest("pointfoot_rough", state, "cuda", seed=9)
        e.split_tail = split
        for k in range(4):
            e.step(state["actions"] * (1.0 - 0.2 * k))
        envs.append(e)
    n = int(envs[0].t["reset_count"])
    assert n > 0 and n == int(envs[1].t["reset_count"])
    for name in ("obs_buf", "priv_obs_buf", "rew_buf", "reset_buf", "episode_sums", "episode_stats", "root_states",
                 "terrain_levels", "extras_time_outs"):
        if name in envs[0].t and envs[0].t[name] is not None:
            assert torch.equal(envs[0].t[name], envs[1].t[name]), name
    assert torch.equal(envs[0].t["reset_ids"][:n], envs[1].t["reset_ids"][:n])
    for e in envs:
        e.close()


def test_sharded_equals_unsharded():
    """env_id_offset: RNG draws are functions of the global env id, so two half shards == one full run."""
    import common
    from pointfoot_b200 import synth
    N = 512
    full = synth.make_state(N, "pointfoot_rough", seed=3)
    e = parity.EnvUnderTest("pointfoot_rough", full, "cuda", seed=3)
    a = full["actions"]
    e.step(a)
    halves = []
    for s in range(2):
        sl = slice(s * N // 2, (s + 1) * N // 2)
        part = {}
        for k, v in full.items():
            if not torch.is_tensor(v) or k in ("height_samples", "terrain_origins"):
                part[k] = v
            elif k == "episode_sums":
                part[k] = v[:, sl].contiguous()
            else:
                rows = v.shape[0] // N
                part[k] = v[sl.start * rows:sl.stop * rows].contiguous()
        h = parity.EnvUnderTest("pointfoot_rough", part, "cuda", seed=3, env_id_offset=sl.start)
        h.step(a[sl])
        halves.append(h)
    for name in ("obs_buf", "priv_obs_buf", "rew_buf", "reset_buf", "root_states", "dof_state", "terrain_levels"):
        got = torch.cat([h.t[name] for h in halves])
        assert torch.equal(got, e.t[name]), name
    for h in halves + [e]:
        h.close()


@pytest.mark.parametrize("pinned,split", [(True, False), (False, False), (True, True)])
def test_host_buffer_step_equals_device_step(pinned, split):
    """pf_step_host (PhysX tensors in host memory, PF:41,50): pinned buffers go through the zero-copy gather / scatter
    kernels, pageable ones through cudaMemcpy; both must reproduce the device-resident step bit for bit.  ``split``:
    the pf_step_host_begin / pf_step_host_join pair (results' transfer on the context's side stream), two steps."""
    import ctypes as C
    from pointfoot_b200 import capi, synth
    N = 1234
    state = synth.make_state(N, "pointfoot_rough", seed=11)
    dev = parity.EnvUnderTest("pointfoot_rough", state, "cuda", seed=11)
    host = parity.EnvUnderTest("pointfoot_rough", state, "cuda", seed=11)
    a = state["actions"].contiguous()
    dev.step(a)
    mk = (lambda t: t.contiguous().pin_memory()) if pinned else (lambda t: t.contiguous().clone())
    hs_t = {k: mk(state[k]) for k in ("root_states", "dof_state", "contact_forces", "rigid_body_states")}
    for k in hs_t:                       # the device copies must come from the host buffers: scramble them first
        host.t[k].fill_(7.0) if k != "rigid_body_states" else None
    hs = capi.PfHostState(**{k: v.data_ptr() for k, v in hs_t.items()})
    P = host.P
    act = mk(a)
    obs, priv = mk(torch.zeros(N, P.num_obs)), mk(torch.zeros(N, P.num_priv_obs))
    rew, rst, tq = mk(torch.zeros(N)), mk(torch.zeros(N, dtype=torch.uint8)), mk(torch.zeros(N, P.num_dof))
    host.step_counter += 1
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if split:
        for k in hs_t:                   # second device step from the same synthetic PhysX tensors the host side re-sends
            dev.t[k].copy_(state[k])
        dev.step(a)
        for it in range(2):
            capi.check(host.lib.pf_step_host_begin(host.ctx, C.byref(host.b), C.byref(hs), C.c_void_p(act.data_ptr()), C.c_void_p(obs.data_ptr()),
                                                   C.c_void_p(priv.data_ptr()), C.c_void_p(rew.data_ptr()), C.c_void_p(rst.data_ptr()),
                                                   C.c_void_p(tq.data_ptr()), C.c_int64(host.step_counter + it), st), "pf_step_host_begin")
            capi.check(host.lib.pf_step_host_join(host.ctx, st), "pf_step_host_join")
        host.step_counter += 1
        torch.cuda.synchronize()
    else:
      capi.check(host.lib.pf_step_host(host.ctx, C.byref(host.b), C.byref(hs), C.c_void_p(act.data_ptr()), C.c_void_p(obs.data_ptr()),
                                     C.c_void_p(priv.data_ptr()), C.c_void_p(rew.data_ptr()), C.c_void_p(rst.data_ptr()),
                                     C.c_void_p(tq.data_ptr()), C.c_int64(host.step_counter), st), "pf_step_host")
    assert torch.equal(obs, dev.t["obs_buf"].cpu()) and torch.equal(priv, dev.t["priv_obs_buf"].cpu())
    assert torch.equal(rew, dev.t["rew_buf"].cpu()) and torch.equal(rst.bool(), dev.t["reset_buf"].cpu())
    assert torch.equal(tq, dev.t["torques"].cpu())
    for k in ("root_states", "dof_state", "episode_sums", "reset_ids"):
        assert torch.equal(host.t[k], dev.t[k]), k
    dev.close(); host.close()


# ------------------------------------------------------------------------------------------------ optional branches
@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
@pytest.mark.parametrize("n", [33, 4096])
def test_cuda_all_24_reward_terms_positive_clip_and_termination_add(task, n):
    """PF:1153 dof_vel, :1170 termination, :1174 dof_pos_limits, :1180 dof_vel_limits, :1226 stumble enabled together
    with only_positive_rewards (PF:312-313) and the post-clip termination add (PF:315-318): the table-driven statement
    of the step (PF_MASK_TABLE) against the oracle; pinned to the unmodified reference by the *_allterms golden files."""
    import common
    out = parity.run(task, n, steps=3, mutate=common.all_reward_terms, num_reward_terms=24)
    assert n < 4096 or out[-1]["num_resets"] > 0


def test_all_terms_generic_statement_matches_too(monkeypatch):
    import common
    monkeypatch.setenv("PF_GENERIC_REWARDS", "1")
    parity.run("pointfoot_rough", 1000, steps=2, mutate=common.all_reward_terms, num_reward_terms=24)


@pytest.mark.parametrize("task", ["pointfoot_flat", "pointfoot_rough"])
def test_cuda_push_step_matches_oracle(task):
    """_push_robots (PF:590-611) on the step where common_step_counter % 350 == 0: pf_push_kernel against the oracle
    (bit-exact: quat_rotate is the same op sequence); the *_push349 golden files pin the same to the reference."""
    out = parity.run(task, 5000, steps=3, start_step=348)
    assert out[1]["push_forces"] == 0.0


@pytest.mark.parametrize("mut", ["control_v", "control_t"])
@pytest.mark.parametrize("n", [31, 4096])
def test_cuda_control_types_v_and_t(mut, n):
    """PF:540-545: velocity-target and direct-torque control laws (both the paired and the scalar torque kernel run:
    4096 * 6 is even, the 64-bit path; the scalar kernel is taken for unaligned tensors, see below)"""
    import common
    parity.run("pointfoot_flat", n, steps=3, mutate=getattr(common, mut))


def test_scalar_torque_kernel_control_v_unaligned():
    """pf_pd_torques_kernel (one thread per DOF) is what runs when a tensor is only 4-byte aligned: same bits."""
    import common
    from pointfoot_b200 import synth
    state = synth.make_state(515, "pointfoot_flat", seed=21)
    a = parity.EnvUnderTest("pointfoot_flat", state, "cuda", common.control_v, seed=21)
    b = parity.EnvUnderTest("pointfoot_flat", state, "cuda", common.control_v, seed=21)
    odd = torch.zeros(b.t["torques"].numel() + 1, device="cuda")[1:].view_as(b.t["torques"])
    b.t["torques"] = odd
    b.b.torques = odd.data_ptr()
    for k in range(2):
        a.step(state["actions"] * (1.0 + k)); b.step(state["actions"] * (1.0 + k))
    assert torch.equal(a.t["torques"], b.t["torques"]) and torch.equal(a.t["rew_buf"], b.t["rew_buf"])
    a.close(); b.close()


def test_reciprocal_index_division_is_what_cuda_eager_does():
    """PF:1102 ``(points / horizontal_scale).long()``: ATen's CPU kernel divides, its CUDA kernel multiplies by the
    fp32 reciprocal of the python scalar.  PF_DIV_RECIPROCAL must reproduce the reference run with its tensors on THIS
    device (the oracle-of-record for the index path on a GPU box), PF_DIV_TRUE the CPU run; the two
    differ for a few samples per million, which is counted here, not hidden."""
    import common
    from oracle import pointfoot_ref
    from pointfoot_b200 import capi, synth
    N = 20000
    state = synth.make_state(N, "pointfoot_rough", seed=17)
    want = {}
    for dev in ("cpu", "cuda"):
        orc, _ = common.oracle_env("pointfoot_rough", state, seed=17)
        want[dev] = pointfoot_ref.heights_native_eager(orc, dev).cpu()
    got = {}
    for mode in (capi.DIV_TRUE, capi.DIV_RECIPROCAL):
        e = parity.EnvUnderTest("pointfoot_rough", state, "cuda", seed=17, index_div_mode=mode)
        e.step(state["actions"])
        got[mode] = e.t["measured_heights"].cpu().clone()
        e.close()
    assert torch.equal(got[capi.DIV_TRUE], want["cpu"]), "PF_DIV_TRUE differs from the CPU-eager reference"
    assert torch.equal(got[capi.DIV_RECIPROCAL], want["cuda"]), "PF_DIV_RECIPROCAL differs from the CUDA-eager reference"
    n_diff = int((want["cpu"] != want["cuda"]).sum())
    print(f"CPU-eager vs CUDA-eager height samples that differ: {n_diff} of {want['cpu'].numel()}")
    # the oracle's own statement of the reciprocal mode agrees with the device run as well
    orc, _ = common.oracle_env("pointfoot_rough", state, seed=17, index_div_mode=capi.DIV_RECIPROCAL)
    assert torch.equal(orc.get_heights(), want["cuda"])


def test_pointfoot_class_pushes_without_a_host_sync_and_inside_a_graph():
    """env.py PointFoot._push_robots on step 350 through the public class (eager and GraphedStep) == the oracle"""
    import common
    from pointfoot_b200 import synth
    from pointfoot_b200.env import PointFoot, GraphedStep
    N, task = 300, "pointfoot_flat"
    state = synth.make_state(N, task, seed=4)
    cfg = common.make_cfg(task, N)
    orc, _ = common.oracle_env(task, state, seed=1)
    orc.common_step_counter = 348
    env = PointFoot(cfg, None, None, "cuda:0", True, state=common.clone_state(state), seed=1)
    env.common_step_counter = 348
    env._step_counter.fill_(348)
    a = torch.zeros(N, 6, device="cuda")
    step = GraphedStep(env, a)
    for k in ("episode_length_buf", "last_actions", "last_dof_vel", "feet_air_time", "last_feet_air_time", "contact_filt",
              "first_contact", "last_contacts"):
        getattr(env, k).copy_(state[k].view(getattr(env, k).shape))
    for k in ("root_states", "dof_state"):
        getattr(env, k).copy_(state[k].view(getattr(env, k).shape))
    env.common_step_counter = 348
    env._step_counter.fill_(348)
    for i in range(3):
        step()
        orc.step(torch.zeros(N, 6))
        if i == 0:
            assert float(env.rigid_body_external_forces.abs().max()) == 0.0
    assert env.common_step_counter == 351 == orc.common_step_counter
    assert torch.equal(env.rigid_body_external_forces.cpu(), orc.push_forces)
    assert float(orc.push_forces.abs().max()) > 0
    env.close()


def _actor_heights(cfg):
    """the actor sees the height columns as well (terrain.measure_heights_actor, PF:367-369)"""
    cfg.terrain.measure_heights_actor = True
    cfg.env.num_propriceptive_obs = 27 + 121


@pytest.mark.parametrize("N,grid187,actor", [(777, False, False), (777, True, False), (70001, False, False), (70001, True, False),
                                             (5000, False, True)])
def test_step_outputs_do_not_depend_on_materialising_measured_heights(N, grid187, actor, monkeypatch):
    """measured_heights is written only for a caller that asks (PointFoot(keep_measured_heights=True)); observations,
    rewards and the episode sums -- its readers on the step path -- are bit-identical either way.  Without it the rough
    cfgs take the specialised scan kernel (pf_height_scan_std_kernel: 121 points = 4 passes, 187 = 6; one env per warp at
    777 envs, 8 at 70,001), with it the general one; PF_SCAN_NO_STD=1 (third env) is the general one without the buffer"""
    import common
    from pointfoot_b200 import synth
    from pointfoot_b200.env import PointFoot
    task = "pointfoot_rough"
    state = synth.make_state(N, task, seed=6, adversarial=True)
    envs = []
    for keep, no_std in ((False, False), (True, False), (False, True)):
        cfg = common.make_cfg(task, N, common.grid187 if grid187 else (_actor_heights if actor else None))
        envs.append(PointFoot(cfg, None, None, "cuda:0", True, state=common.clone_state(state), seed=1, keep_measured_heights=keep))
        envs[-1]._no_std = no_std
    real_step = PointFoot.step

    def step(e, a):                                    # the switch is read at every launch
        if e._no_std:
            monkeypatch.setenv("PF_SCAN_NO_STD", "1")
        else:
            monkeypatch.delenv("PF_SCAN_NO_STD", raising=False)
        out = real_step(e, a)
        torch.cuda.synchronize()
        return out
    a = synth.make_state(N, task, seed=7)["actions"].cuda()
    for _ in range(3):
        outs = [step(e, a) for e in envs]
        for other in (1, 2):
            for x, y in zip(outs[0][:4], outs[other][:4]):
                assert torch.equal(x, y)
            assert torch.equal(envs[0]._episode_sums, envs[other]._episode_sums)
    monkeypatch.delenv("PF_SCAN_NO_STD", raising=False)
    assert envs[1].measured_heights.shape == (N, 187 if grid187 else 121) and float(envs[1].measured_heights.abs().sum()) > 0
    with pytest.raises(AttributeError, match="keep_measured_heights"):
        envs[0].measured_heights
    for e in envs:
        e.close()
