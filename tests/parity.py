"""Parity driver shared by the GPU tests, ``__graft_entry__.smoke()`` and bench.py's self-check: runs the env
side of the hot path through the C-ABI and compares every output with the oracle (oracle/pointfoot_ref.py) on
the same seeded synthetic state.  TEST INFRASTRUCTURE.

Bars (BASELINE.json north_star): integer / boolean / index outputs bit-exact; fp32 outputs within
RTOL/ATOL below (most of them are in fact bit-identical: only expf / atan2f and the order of the 121-term
base-height mean differ between CUDA and ATen CPU).
"""
import ctypes as C

import torch

import common
from pointfoot_b200 import capi, synth

RTOL, ATOL = 1e-5, 2e-6

EXACT = ("reset_buf", "time_out_buf", "terminate", "episode_length_buf", "contact_filt", "first_contact",
         "last_contacts", "terrain_levels", "extras_time_outs")
# fp32 state that is bit-identical by construction (no libm, no long sums)
EXACT_F32 = ("torques", "actions", "base_lin_vel", "base_ang_vel", "projected_gravity", "last_actions",
             "last_dof_vel", "last_root_vel", "feet_air_time", "last_feet_air_time", "current_max_feet_height",
             "last_max_feet_height", "root_states", "dof_state", "env_origins", "measured_heights", "push_forces")
CLOSE = ("obs_buf", "priv_obs_buf", "rew_buf", "episode_sums", "commands")


def oracle_field(orc, name):
    if name == "measured_heights":
        return orc.measured_heights if torch.is_tensor(orc.measured_heights) else None
    return getattr(orc, name, None)


class EnvUnderTest:
    """pf_ctx + buffers for one synthetic state, stepped through the C-ABI (or the host emulation)."""

    def __init__(self, task, state, device, mutate=None, backend="cuda", **kw):
        self.cfg, self.P, self.noise, self.pts, self.names = common.make_params(task, state, mutate, **kw)
        self.t, self.b = common.alloc_buffers(self.P, state, device)
        self.t["push_forces"] = torch.zeros(self.P.num_envs, self.P.num_bodies, 3, device=device)
        self.t["base_mass"] = state["base_mass"].clone().to(device)
        self.push_interval, self.sim_dt_py, self.max_push_vel_xy_py = common.push_schedule(self.cfg)
        self.backend = backend
        self.step_counter = int(state.get("common_step_counter", 0))
        self.device = device
        if backend == "cuda":
            self.lib = capi.load()
            self.ctx = C.c_void_p()
            noise, pts = self.noise.contiguous(), self.pts.contiguous()
            capi.check(self.lib.pf_ctx_create(C.byref(self.P), C.c_void_p(noise.data_ptr()),
                                              C.c_void_p(pts.data_ptr() if pts.numel() else None), C.byref(self.ctx)),
                       "pf_ctx_create")
        else:
            self.lib = common.host_emu()

    def close(self):
        if self.backend == "cuda" and self.ctx:
            self.lib.pf_ctx_destroy(self.ctx)
            self.ctx = None

    def step(self, actions, physics=None):
        a = actions.to(self.device).contiguous()
        self.step_counter += 1
        if self.backend == "cuda":
            stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            for i in range(self.P.decimation):
                capi.check(self.lib.pf_pd_torques(self.ctx, C.byref(self.b), C.c_void_p(a.data_ptr() if i == 0 else None),
                                                  stream), "pf_pd_torques")
                if physics is not None:
                    physics(self)
            if self.push_interval and self.step_counter % self.push_interval == 0:      # PF:492-493, 590-611
                m = float(self.t["base_mass"].cpu().mean().item()) * self.max_push_vel_xy_py / self.sim_dt_py
                capi.check(self.lib.pf_push_forces(self.ctx, C.byref(self.b), C.c_void_p(self.t["push_forces"].data_ptr()),
                                                   C.c_float(m), C.c_int64(self.step_counter), stream), "pf_push_forces")
            if getattr(self, "split_tail", False):     # tail on the context's side stream, other work queued beside it
                capi.check(self.lib.pf_post_physics_step_begin(self.ctx, C.byref(self.b), C.c_int64(self.step_counter), stream),
                           "pf_post_physics_step_begin")
                self.t["obs_buf"].sum()
                capi.check(self.lib.pf_post_physics_step_join(self.ctx, stream), "pf_post_physics_step_join")
            else:
                capi.check(self.lib.pf_post_physics_step(self.ctx, C.byref(self.b), C.c_int64(self.step_counter), stream),
                           "pf_post_physics_step")
            torch.cuda.synchronize()
        else:
            for i in range(self.P.decimation):
                self.lib.pf_emu_pd_torques(C.byref(self.P), C.byref(self.b), C.c_void_p(a.data_ptr() if i == 0 else None))
                if physics is not None:
                    physics(self)
            if self.push_interval and self.step_counter % self.push_interval == 0:
                m = float(self.t["base_mass"].cpu().mean().item()) * self.max_push_vel_xy_py / self.sim_dt_py
                self.lib.pf_emu_push_forces(C.byref(self.P), C.byref(self.b), C.c_void_p(self.t["push_forces"].data_ptr()),
                                            C.c_float(m), C.c_longlong(self.step_counter))
            self.lib.pf_emu_post_physics_step(C.byref(self.P), C.byref(self.b), C.c_void_p(self.noise.data_ptr()),
                                              C.c_void_p(self.pts.data_ptr()), C.c_longlong(self.step_counter))


def compare(orc, env, label=""):
    """Raises AssertionError naming the first field that breaks its bar; returns a dict of max abs diffs."""
    t = env.t
    diffs = {}

    def get(name):
        o = oracle_field(orc, name)
        g = t.get(name)
        if o is None or g is None:
            return None, None
        return o, g.cpu().view(o.shape)

    for name in EXACT + EXACT_F32:
        o, g = get(name)
        if o is None:
            continue
        if not torch.equal(o, g.to(o.dtype)):
            bad = (o != g.to(o.dtype)).nonzero()
            raise AssertionError(f"{label}{name}: not bit-exact at {bad[:4].tolist()} ({bad.shape[0]} mismatches)")
        diffs[name] = 0.0
    for name in CLOSE:
        o, g = get(name)
        if o is None:
            continue
        torch.testing.assert_close(g, o, rtol=RTOL, atol=ATOL, msg=lambda m, n=name: f"{label}{n}: {m}")
        diffs[name] = (g - o).abs().max().item()
    n = int(t["reset_count"].item())
    ids = t["reset_ids"][:n].cpu()
    if not torch.equal(ids, orc.reset_ids):
        raise AssertionError(f"{label}reset_ids differ: {ids.tolist()[:8]} vs {orc.reset_ids.tolist()[:8]}")
    if n:
        torch.testing.assert_close(t["episode_stats"].cpu(), orc.episode_stats, rtol=1e-5, atol=1e-6,
                                   msg=lambda m: f"{label}episode_stats: {m}")
    diffs["num_resets"] = n
    return diffs


def run(task, num_envs, steps=3, device="cuda", backend="cuda", adversarial=True, mutate=None, seed=1,
        feeder=None, start_step=0, num_reward_terms=19, **kw):
    """Step oracle and implementation side by side; ``feeder(i) -> dict of PhysX tensors`` replaces the
    physics between steps (default: state tensors stay, apart from the reset rows both sides write)."""
    state = synth.make_state(num_envs, task, seed=seed, adversarial=adversarial, num_reward_terms=num_reward_terms)
    state["common_step_counter"] = start_step
    orc, _ = common.oracle_env(task, state, mutate, seed=seed, **kw)
    env = EnvUnderTest(task, state, device, mutate, backend, seed=seed, **kw)
    out = []
    try:
        for i in range(steps):
            if feeder is not None:
                new = feeder(i)
                for k, v in new.items():
                    getattr(orc, k).view(v.shape).copy_(v)
                    env.t[k].view(v.shape).copy_(v.to(device))
            a = synth.make_state(num_envs, task, seed=seed + 100 + i)["actions"] * (30.0 if i == 1 else 1.0)
            orc.step(a.clone())
            env.step(a.clone())
            out.append(compare(orc, env, f"[{task} N={num_envs} step {i}] "))
    finally:
        env.close()
    return out


def smoke_check():
    d = run("pointfoot_rough", 64, steps=2)
    print("env parity:", d[-1])
    try:
        import parity_rl
    except ImportError:
        return
    parity_rl.smoke_check()


# ------------------------------------------------------------------------------------------------ golden
import hashlib
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# file -> (task, cfg mutation, common_step_counter the run starts from).  All produced by tests/golden/make_golden.py from the
# UNMODIFIED reference point_foot.py.
GOLDEN = {"pointfoot_flat_n64.npz": ("pointfoot_flat", None, 0),
          "pointfoot_rough_n64.npz": ("pointfoot_rough", None, 0),
          "pointfoot_rough_p187_n64.npz": ("pointfoot_rough", common.grid187, 0),
          # _push_robots fires on the first step (350 % 350 == 0, PF:492-493, 590-611)
          "pointfoot_flat_push349_n64.npz": ("pointfoot_flat", None, 349),
          "pointfoot_rough_push349_n64.npz": ("pointfoot_rough", None, 349),
          # the five reward terms the pointfoot cfgs disable + only_positive_rewards + the termination add (PF:312-318)
          "pointfoot_rough_allterms_n64.npz": ("pointfoot_rough", common.all_reward_terms, 0),
          "pointfoot_flat_allterms_n64.npz": ("pointfoot_flat", common.all_reward_terms, 0),
          # control_type 'V' / 'T' (PF:540-545)
          "pointfoot_flat_ctrlV_n64.npz": ("pointfoot_flat", common.control_v, 0),
          "pointfoot_flat_ctrlT_n64.npz": ("pointfoot_flat", common.control_t, 0)}
_GOLD_EXACT = ("reset", "terminate", "time_out_buf", "episode_length_buf", "contact_filt", "first_contact",
               "last_contacts", "terrain_levels", "time_outs", "torques", "base_lin_vel", "base_ang_vel", "projected_gravity",
               "last_actions", "last_dof_vel", "last_root_vel", "feet_air_time", "last_feet_air_time",
               "current_max_feet_height", "last_max_feet_height", "root_states", "dof_state", "env_origins",
               "measured_heights", "push_forces")
_GOLD_CLOSE = ("obs", "priv", "rew", "episode_sums", "commands")
_GOLD_NAME = {"time_outs": "extras_time_outs", "reset": "reset_buf", "obs": "obs_buf", "priv": "priv_obs_buf", "rew": "rew_buf"}


def load_golden(fname):
    """-> (task, mutate, state dict of torch tensors, npz)"""
    task, mutate, start_step = GOLDEN[fname]
    z = np.load(os.path.join(GOLDEN_DIR, fname))
    state = {k[3:]: torch.from_numpy(z[k].copy()) for k in z.files if k.startswith("in_")}
    state["common_step_counter"] = start_step
    assert int(z["start_step"]) == start_step if "start_step" in z.files else start_step == 0
    if task == "pointfoot_rough":
        hf, origins = synth.make_heightfield(1)
        sha = hashlib.sha256(hf.numpy().tobytes()).hexdigest()
        assert sha == bytes(z["hf_sha256"]).decode(), "synthetic heightfield drifted from the one the golden run used"
        state["height_samples"] = hf
    return task, mutate, state, z


def golden_actions(z, i):
    task_n = z["in_actions"].shape[0]
    return None, task_n


def check_against_golden(fname, stepper, fetch, steps=3):
    """``stepper(i, actions)`` advances the implementation; ``fetch(name)`` returns a CPU tensor (or None).
    The reference keeps ``episode_sums`` / ``extras["episode"]`` as dicts in ``dir()`` order with 'termination' at its
    alphabetical place (PF:817-820); the C-ABI's [R,N] rows are the summation order with 'termination' last
    (include/pointfoot_b200.h), so rows are matched BY NAME."""
    task, mutate, state, z = load_golden(fname)
    N = state["root_states"].shape[0]
    ref_names = [str(n) for n in z["reward_names"]]
    rows = [ref_names.index(n) for n in fetch("reward_names")]
    assert len(rows) == len(ref_names), "enabled reward terms differ from the reference's"
    for i in range(steps):
        a = synth.make_state(N, task, seed=1 + 100 + i)["actions"] * (30.0 if i == 1 else 1.0)
        stepper(i, a)
        p = f"s{i}_"
        for name in _GOLD_EXACT + _GOLD_CLOSE:
            if p + name not in z.files:
                continue
            got = fetch(_GOLD_NAME.get(name, name))
            if got is None:
                continue
            want = torch.from_numpy(z[p + name])
            if name == "episode_sums":
                want = want[rows]
            got = got.cpu().reshape(want.shape).to(want.dtype)
            if name in _GOLD_EXACT:
                assert torch.equal(got, want), f"[{fname} step {i}] {name} not bit-exact vs the reference"
            else:
                torch.testing.assert_close(got, want, rtol=RTOL, atol=ATOL,
                                           msg=lambda m, n=name: f"[{fname} step {i}] {n}: {m}")
        ids = fetch("reset_ids")
        assert torch.equal(ids.cpu().to(torch.int32), torch.from_numpy(z[p + "reset_ids"])), f"[{fname} step {i}] reset ids"
        want = z[p + "extras_episode"]
        if want.size:
            stats = fetch("episode_stats")[:want.size].cpu()
            torch.testing.assert_close(stats, torch.from_numpy(want)[rows], rtol=1e-5, atol=1e-6)


def golden_oracle(fname):
    """the CPU restatement against the unmodified reference's outputs"""
    task, mutate, state, _ = load_golden(fname)
    orc, _ = common.oracle_env(task, state, mutate)

    def fetch(name):
        if name == "reset_ids":
            return orc.reset_ids
        if name == "reward_names":
            return orc.reward_names + (["termination"] if orc.P.has_termination_term else [])
        return oracle_field(orc, name)
    check_against_golden(fname, lambda i, a: orc.step(a.clone()), fetch)


def golden_impl(fname, device="cuda", backend="cuda"):
    """the CUDA path (or the host emulation of its per-env logic) against the unmodified reference's outputs"""
    task, mutate, state, _ = load_golden(fname)
    env = EnvUnderTest(task, state, device, mutate, backend)

    def fetch(name):
        if name == "reset_ids":
            return env.t["reset_ids"][:int(env.t["reset_count"].item())]
        if name == "reward_names":
            return list(env.names)
        return env.t.get(name)
    try:
        check_against_golden(fname, lambda i, a: env.step(a.clone()), fetch)
    finally:
        env.close()
