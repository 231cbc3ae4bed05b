"""Headline benchmark: env-steps/s of the fused step (4x PD torques + post_physics_step + ActorCritic act/evaluate) on
synthetic state tensors.  Contract: ``python bench.py --gpus N --steps K --warmup W`` prints ONE JSON line (rank 0).

A "step" is one pass of the hot path over one batch of N envs per GPU: clip actions, decimation x _compute_torques,
post_physics_step (point_foot.py:170-230) and the ActorCritic act/evaluate that produces the next actions.

What the line holds (BASELINE.json: "env-steps/sec (step+policy) at 4096-262k envs, 1/2/4/8 B200; % HBM roofline"):
  value / ms_per_step   configs[1]: pointfoot_flat, 4,096 envs per GPU, inputs resident in HBM, one CUDA graph per step
  value_fp32            the same with the fp32 SIMT policy instead of the bf16 tcgen05 one
  e2e                   the same metric through the host-buffer C-ABI (pf_step_host_begin/_join): PhysX tensors + actions
                        H2D, results D2H, every step
  roofline              pf_post_physics_step alone against the measured HBM peak (MEASURED_PEAKS.json)
  sweep                 flat / rough P=121 / rough P=187 x {4,096, 65,536, 262,144} envs: ms per step, env-steps/s, e2e,
                        roofline fractions of the env step and of the act kernel (1 GPU); on N GPUs: config C5 -- 262,144
                        envs sharded N ways (strong) and 32,768 envs per GPU (weak)
  ppo_iteration         configs[3]: 24 x 4,096 rollout bookkeeping + GAE + 5 x 4 minibatch updates, fp32 and bf16, flat and
                        rough nets; on N GPUs with the NCCL gradient / statistics all-reduces (their cost reported)
  cuda_eager_baseline   the UNMODIFIED reference point_foot.py (baseline/_ref) with its tensors on this GPU, torch eager
  cpu_baseline          the same reference on the host cores

``--impl reference`` times the reference's CPU implementation of the path: the unmodified ``point_foot.py`` installed in
baseline/_ref (baseline/install_reference.py) behind the stubs of oracle/ref_harness.py, with the restated rsl_rl
ActorCritic (oracle/rsl_rl_ref.py: the package itself is not installable offline); if baseline/_ref is absent, the oracle
port.  Physics (gym.simulate) is excluded on both sides: Isaac Gym cannot be installed.
"""
import argparse
import ctypes as C
import hashlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "env_steps_per_sec"
UNIT = "env-steps/s"
TASKS = {"pointfoot_flat": ("PointFootFlatCfg", "PointFootFlatCfgPPO"),
         "pointfoot_rough": ("PointFootRoughCfg", "PointFootRoughCfgPPO")}
MIN_TIMED_S = 0.5            # the K-step loop is repeated until this much DEVICE time has been measured


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--task", default="pointfoot_flat", choices=sorted(TASKS))
    p.add_argument("--num-envs", type=int, default=4096, help="envs PER GPU (weak scaling)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--no-policy", action="store_true")
    p.add_argument("--no-sweep", action="store_true", help="headline only (profiling runs)")
    p.add_argument("--no-ppo", action="store_true")
    p.add_argument("--no-eager", action="store_true")
    p.add_argument("--min-seconds", type=float, default=MIN_TIMED_S)
    p.add_argument("--l2", default="flush", choices=["flush", "none"],
                   help="'none' is a DIAGNOSTIC (hot caches): such a line is marked invalid")
    return p.parse_args()


def make_cfgs(task, num_envs, grid187=False):
    from pointfoot_b200 import config
    cfg = getattr(config, TASKS[task][0])()
    train_cfg = getattr(config, TASKS[task][1])()
    cfg.env.num_envs = num_envs
    if grid187:             # the 17 x 11 LeggedRobotCfg scan grid (legged_robot_config.py:54-55)
        cfg.terrain.measured_points_x = [round(-0.8 + 0.1 * i, 1) for i in range(17)]
        cfg.terrain.measured_points_y = [round(-0.5 + 0.1 * i, 1) for i in range(11)]
        cfg.env.num_privileged_obs = 27 + 187
    return cfg, train_cfg


def workload_name(task, num_envs, decimation=4, policy=True):
    """identical in both arms (the driver compares the strings)"""
    return (f"{task} {num_envs} envs: fused PD torques x{decimation} + post_physics_step + reward/obs"
            + (" + ActorCritic act/evaluate" if policy else ""))


def bench_config(task, n_per_gpu, world, policy=True):
    """``config`` of the JSON line: the same dict in the b200 and the reference arm."""
    return {"workload": workload_name(task, n_per_gpu * world, 4, policy), "task": task, "num_envs": n_per_gpu * world,
            "num_envs_per_gpu": n_per_gpu, "policy": "ActorCritic [128,64,32] ELU (flat) / [512,256,128] (rough), act + evaluate + "
            "sample + log-prob" if policy else "none",
            "physics": "excluded (Isaac Gym not installable): synthetic state tensors",
            "terrain": "rough task: int16[1300,2100] field generated on the GPU by pf_terrain_generate (reference curriculum layout, "
                       "10 x 20 sub-terrains) before any timed region; flat task: plane",
            "parallelism": f"env-sharded x{world} (rollout: no data-path collective; PPO.update: NCCL all-reduce of the flat "
                           "gradient + loss statistics per optimizer step, see ppo_iteration)"}


def algorithmic_bytes_per_env(P):
    """compulsory fp32 traffic of one policy step, per env.
    post_physics_step: READ root 52 + dof_state 48 + contact 108 + feet xyz 24 + actions 24 + torques 24 +
    last_actions 24 + last_dof_vel 24 + commands 4C + ep_len 8 + feet state machine 38 + episode_sums 4R;
    WRITE obs 108 + priv 4*PO + rew 4 + flags 3 + base vel/gravity 36 + last_* 72 + ep_len 8 + feet SM 46 +
    episode_sums 4R.  Torques: 4 x (actions 24 + dof_state 48 + torques 24).
    PPO.act: READ obs 4*O + critic obs 4*PO; WRITE the rollout-storage row (obs 4*O, critic obs 4*PO, actions / mu / sigma
    3 x 4*A, value + log-prob 8) + the actions handed to the env 4*A."""
    C_, R = P.num_commands, P.num_reward_terms + P.has_termination_term
    po = P.num_priv_obs or P.num_obs
    read = 52 + 48 + 108 + 24 + 24 + 24 + 24 + 24 + 4 * C_ + 8 + 38 + 4 * R
    write = 4 * P.num_obs + 4 * P.num_priv_obs + 4 + 3 + 36 + 72 + 8 + 46 + 4 * R
    act = 4 * P.num_obs + 4 * po + (4 * P.num_obs + 4 * po + 3 * 4 * P.num_dof + 8) + 4 * P.num_dof
    return {"post_physics": read + write, "torques": P.decimation * 96, "act": act}


def mlp_flops_per_row(dims_a, dims_c):
    return 2 * (sum(a * b for a, b in zip(dims_a[:-1], dims_a[1:])) + sum(a * b for a, b in zip(dims_c[:-1], dims_c[1:])))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.samples, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ reference legs
def reference_root():
    """the offline install of the unmodified reference that travels with the snapshot; never /root/reference at run time"""
    root = os.path.join(ROOT, "baseline", "_ref")
    return root if os.path.isfile(os.path.join(root, "legged_gym/envs/pointfoot/point_foot.py")) else None


class ReferenceStep:
    """One policy step of the reference stack on ``device``: the unmodified PointFoot.step (4 x _compute_torques,
    post_physics_step; gym.* inert) + rsl_rl ActorCritic act / evaluate / log-prob (restated: oracle/rsl_rl_ref.py).
    Falls back to the oracle port of the env (CPU only) when baseline/_ref is missing."""

    def __init__(self, task, num_envs, device="cpu", with_policy=True):
        import torch
        from pointfoot_b200 import synth
        self.torch, self.device, self.num_envs = torch, device, num_envs
        state = synth.make_state(num_envs, task, seed=1)
        root = reference_root()
        self.kind = "reference" if root else "port"
        if root:
            os.environ["POINTFOOT_REFERENCE_ROOT"] = root
            from oracle import ref_harness
            assert ref_harness.REFERENCE_ROOT == root, "oracle.ref_harness was imported before the root was set"
            self.env = ref_harness.build_reference_env(task, state, device=device, contract_rng=False)
            self.step_env = lambda a: self.env.step(a)[:2]
            no, npv, na = self.env.num_obs, self.env.num_privileged_obs, self.env.num_actions
        else:
            if device != "cpu":
                raise RuntimeError("baseline/_ref is missing: the CUDA-eager leg needs the reference package")
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import common
            self.env, (cfg, P, noise, pts, names) = common.oracle_env(task, state)
            self.step_env = lambda a: self.env.step(a)[:2]
            no, npv, na = P.num_obs, P.num_priv_obs or None, P.num_dof
        self.policy = None
        if with_policy:
            from oracle import rsl_rl_ref
            _, train_cfg = make_cfgs(task, num_envs)
            torch.manual_seed(1)
            pol = rsl_rl_ref.ActorCritic(no, npv or no, na, train_cfg.policy.actor_hidden_dims, train_cfg.policy.critic_hidden_dims)
            pol.to(device)
            if device != "cpu":          # the restatement's sampling contract is a numpy statement; on the device draw with torch
                def act(obs, pol=pol):
                    pol.update_distribution(obs)
                    return pol.distribution.sample()
                pol.act = act
            self.policy = pol
        self.actions = state["actions"].clone().to(device)

    def __call__(self):
        torch = self.torch
        obs, priv = self.step_env(self.actions)
        if self.policy is not None:
            with torch.inference_mode():
                self.actions = self.policy.act(obs).detach()
                self.policy.evaluate(priv if priv is not None else obs)
                self.policy.get_actions_log_prob(self.actions)


def cpu_env_steps_per_sec(task, num_envs, budget_s=12.0, min_steps=5, threads=None):
    import torch
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = ReferenceStep(task, num_envs, "cpu")
    for _ in range(3):
        ref()
    n, t0 = 0, time.perf_counter()
    while True:
        ref()
        n += 1
        el = time.perf_counter() - t0
        if (el > budget_s and n >= min_steps) or n >= 100000:
            break
    return num_envs * n / el, n, cores, ref.kind


def cuda_eager_steps_per_sec(task, num_envs, device, flush, budget_s=4.0):
    """the torch-eager bar: the reference's own code with its tensors on this GPU"""
    import torch
    ref = ReferenceStep(task, num_envs, device)
    for _ in range(5):
        ref()
    torch.cuda.synchronize()
    ms, n = [], 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget_s or n < 5:
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); ref(); b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b)); n += 1
    ms.sort()
    med = ms[len(ms) // 2]
    return {"value": num_envs / (med * 1e-3), "unit": UNIT, "ms_per_step": med, "steps": n, "kind": ref.kind,
            "what": f"unmodified legged_gym point_foot.py ({task}, {num_envs} envs) + rsl_rl ActorCritic (restated) as torch "
                    "eager on this GPU, CUDA events, L2 flushed between steps, physics excluded; torch.save of PF:334 skipped"}


def reference_line(args, v, steps, warm, el, cores, kind, n_total):
    cfg = bench_config(args.task, args.num_envs, args.gpus, True)
    what = ("unmodified legged_gym point_foot.py from baseline/_ref behind oracle/ref_harness.py" if kind == "reference"
            else "oracle port of the reference torch code (baseline/_ref missing)")
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1e3 * el / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{steps} steps of {n_total} envs, {what} + rsl_rl ActorCritic (restated), torch CPU eager, "
                                       f"{cores} threads"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path, all host threads, same config."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    steps, warm = max(1, args.steps), max(3, args.warmup)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_total = args.num_envs * args.gpus
    ref = ReferenceStep(args.task, n_total, "cpu")
    for _ in range(warm):
        ref()
    t0 = time.perf_counter()
    for _ in range(steps):
        ref()
    el = time.perf_counter() - t0
    print(json.dumps(reference_line(args, n_total * steps / el, steps, warm, el, cores, ref.kind, n_total)), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
_TERRAIN = {}


def rough_terrain(device):
    """The height field of the rough task: generated ONCE per process on the GPU by pointfoot_b200.terrain.Terrain (the reference's
    curriculum layout, terrain.py:85-92: 10 x 20 sub-terrains of 8 m, 25 m border -> int16[1300, 2100]) -- outside every timed region,
    like Terrain.__init__ in the reference.  Returns (height_samples, terrain_origins fp32[10, 20, 3]) as synth.make_state expects."""
    key = str(device)
    if key not in _TERRAIN:
        import torch
        from pointfoot_b200 import config
        from pointfoot_b200.terrain import Terrain
        t = Terrain(config.PointFootRoughCfg.terrain(), 4096, device=device, seed=1)
        _TERRAIN[key] = (t.heightsamples.cpu(), torch.from_numpy(t.env_origins).to(torch.float))
    return _TERRAIN[key]


class Workload:
    """env + policy + the CUDA-graphed step for one (task, envs per GPU, scan grid) on this rank."""

    def __init__(self, task, num_envs, device, rank, world, grid187=False, precision="bf16", policy=True, graph=True,
                 terrain=None):
        import numpy as np
        import torch
        from pointfoot_b200 import synth
        from pointfoot_b200.env import PointFoot, GraphedStep
        self.task, self.num_envs, self.world = task, num_envs, world
        cfg, train_cfg = make_cfgs(task, num_envs, grid187)
        if task == "pointfoot_rough" and terrain is None:
            terrain = rough_terrain(device)
        state = synth.make_state(num_envs, task, seed=1, env_id_offset=rank * num_envs, terrain=terrain)
        if task == "pointfoot_rough":     # terrain_types is a function of the GLOBAL env id (point_foot.py:981-983)
            gid = np.arange(num_envs) + rank * num_envs
            state["terrain_types"] = torch.from_numpy(np.floor(gid / (num_envs * world / 20)).astype(np.int64))
        self.env = PointFoot(cfg, None, None, device, True, state=state, seed=train_cfg.seed, env_id_offset=rank * num_envs)
        self.cfg, self.train_cfg = cfg, train_cfg
        self.policy = None
        if policy:
            from pointfoot_b200 import rl
            self.policy = rl.make_bench_policy(self.env, train_cfg, precision)
        self.actions = self.policy.actions if self.policy is not None else torch.zeros(num_envs, self.env.num_actions, device=device)
        extra = (lambda: self.policy.act_into(self.env, self.actions)) if self.policy is not None else None
        self.env.reset()
        lib = self.env._lib
        if graph:
            self.step_fn = GraphedStep(self.env, self.actions, extra)
            # decimation x pf_pd_torques, [pf_height_scan_kernel], pf_post_physics_kernel (+ its tail), the act kernel(s):
            # COUNTED by the library's launch counter while the step was captured
            self.launches_per_step = self.step_fn.launches_per_step
        else:
            def step_fn():
                self.env.step(self.actions)
                if extra:
                    extra()
            self.step_fn = step_fn
            n0 = lib.pf_debug_launch_count()
            step_fn()
            self.launches_per_step = int(lib.pf_debug_launch_count() - n0)

    def close(self):
        if self.policy is not None:
            self.policy.alg.close()
        self.env.close()


def time_rounds(step_fn, steps, warmup, flush, barrier, world, device, min_seconds):
    """W untimed warm-up steps, then rounds of EXACTLY K timed steps (each step between its own pair of CUDA events on the
    launching stream, the L2 flush in front of it excluded), barrier + synchronize on both sides of every round, MAX over
    ranks per round.  Rounds are repeated until ``min_seconds`` of device time have been measured, so that a 0.8 ms region
    (20 x 39 us) cannot carry the headline; the reported figure is the MEDIAN round.  Returns (ms per round, rounds, all)."""
    import torch
    import torch.distributed as dist

    def one_round():
        barrier()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in evs:
            flush()
            a.record()
            step_fn()
            b.record()
        torch.cuda.synchronize()
        barrier()
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(warmup):
        flush()
        step_fn()
    rounds = [one_round()]
    want = int(min(2000, max(1, math.ceil(min_seconds * 1e3 / max(rounds[0], 1e-3)))))
    for _ in range(want - 1):
        rounds.append(one_round())
    s = sorted(rounds)
    return s[len(s) // 2], len(rounds), rounds


def time_call(fn, flush, n=20, warm=3):
    """mean device time (ms) of ``fn`` alone: CUDA events, L2 flushed before every call, synchronize on both sides"""
    import torch
    tot = 0.0
    for i in range(n + warm):
        flush()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); e.record()
        torch.cuda.synchronize()
        if i >= warm:
            tot += a.elapsed_time(e)
    return tot / n


STEP_KERNEL_SOURCES = ("pf_env.cu", "pf_env_core.cuh", "pf_rng.cuh", "pf_ctx.h", "pf_launch.h")


def source_hash():
    """hash of the files that define the step kernels (scan, step, finalize): what `roofline.traffic` depends on"""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "pointfoot_b200", "csrc")
    for f in STEP_KERNEL_SOURCES:
        with open(os.path.join(d, f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_traffic(task, num_envs):
    """DRAM bytes per launch of the step kernels from THIS build's ncu --set full capture (profiles/r2_ncu_traffic.json
    records the hash of csrc/ it was taken on); null when the sources have changed since -- never a stale constant."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")) as f:
            t = json.load(f)
        if t.get("source_hash") != source_hash():
            return None
        return t.get(f"{task}:{num_envs}", {}).get("bytes")
    except (OSError, ValueError):
        return None


def rooflines(w, flush, peaks, n=20):
    """pf_post_physics_step alone (scan + step kernel + its tail) and PPO.act alone, each against its bound."""
    import torch
    from pointfoot_b200 import capi
    env = w.env
    lib, ctx, b = env._lib, env._ctx, C.byref(env._bufs)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def post():
        capi.check(lib.pf_post_physics_step(ctx, b, C.c_int64(0), st), "pf_post_physics_step")
    ms = time_call(post, flush, n)
    ab = algorithmic_bytes_per_env(env._P)
    nbytes = ab["post_physics"] * env.num_envs
    hbm = peaks.get("hbm_gbs", 6650.0)
    src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    achieved = nbytes / (ms * 1e-3) / 1e9
    kern = "pf_post_physics_kernel" + (" + pf_height_scan_kernel" if env._P.num_height_points > 0 else "") + env.tail_kernel_name
    post_r = {"bound": "hbm", "kernel": kern, "achieved": achieved, "peak": hbm, "peak_source": src, "unit": "GB/s",
              "frac": achieved / hbm, "traffic": ncu_traffic(w.task, env.num_envs), "us_per_launch": ms * 1e3,
              "algorithmic_bytes_per_launch": nbytes}
    act_r = None
    if w.policy is not None:
        ms_a = time_call(lambda: w.policy.act_into(env, w.actions), flush, n)
        abytes = ab["act"] * env.num_envs
        flops = mlp_flops_per_row(w.policy.ac.actor_dims, w.policy.ac.critic_dims) * env.num_envs
        tf_peak = peaks.get("bf16_tflops_sustained", 1381.4)
        hbm_frac = abytes / (ms_a * 1e-3) / 1e9 / hbm
        tc_frac = flops / (ms_a * 1e-3) / 1e12 / tf_peak
        act_r = {"kernel": w.policy.kernel_name + (" (PPO.act: both nets + sample + log-prob + storage row)" if w.policy.precision == "bf16" else ""),
                 "us_per_launch": ms_a * 1e3,
                 "bound": "tensor" if tc_frac > hbm_frac else "hbm", "hbm_frac": hbm_frac, "tensor_frac": tc_frac,
                 "achieved_gbs": abytes / (ms_a * 1e-3) / 1e9, "achieved_tflops": flops / (ms_a * 1e-3) / 1e12,
                 "peak_tflops": tf_peak, "algorithmic_bytes_per_launch": abytes, "algorithmic_flops_per_launch": flops}
    return post_r, act_r


def e2e_host(w, steps, warmup, barrier, use_graph=True):
    """Same metric through the host-buffer C-ABI call: PhysX state tensors + actions H2D, results D2H, per step."""
    import torch
    from pointfoot_b200 import capi
    env, policy = w.env, w.policy
    P, N = env._P, env.num_envs
    pin = lambda t: t.detach().cpu().contiguous().pin_memory()
    hs_t = {k: pin(getattr(env, k)) for k in ("root_states", "dof_state", "contact_forces", "rigid_body_states")}
    hs = capi.PfHostState(**{k: v.data_ptr() for k, v in hs_t.items()})
    act = pin(torch.zeros(N, P.num_dof))
    obs = pin(torch.zeros(N, P.num_obs)); priv = pin(torch.zeros(N, max(P.num_priv_obs, 1)))
    rew = pin(torch.zeros(N)); rst = pin(torch.zeros(N, dtype=torch.uint8)); tq = pin(torch.zeros(N, P.num_dof))
    lib, ctx, b = env._lib, env._ctx, C.byref(env._bufs)

    def enqueue(sync):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        fn = lib.pf_step_host if sync else lib.pf_step_host_begin
        capi.check(fn(ctx, b, C.byref(hs), C.c_void_p(act.data_ptr()), C.c_void_p(obs.data_ptr()),
                      C.c_void_p(priv.data_ptr()) if P.num_priv_obs else None, C.c_void_p(rew.data_ptr()),
                      C.c_void_p(rst.data_ptr()), C.c_void_p(tq.data_ptr()), C.c_int64(0), st), "pf_step_host")
        if policy is not None:      # same metric as `value`: the policy runs on the device observations and its actions
            policy.act_into(env, policy.actions)          # go back to the host for the next physics step
            act.copy_(policy.actions, non_blocking=True)
        if not sync:                # the results' transfer ran beside the policy kernel; the stream waits for it here
            capi.check(lib.pf_step_host_join(ctx, st), "pf_step_host_join")

    graph = None
    if policy is not None and use_graph:
        for _ in range(2):
            enqueue(False)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            enqueue(False)

    def one():
        env.common_step_counter += 1
        if graph is not None:
            graph.replay()
            torch.cuda.current_stream().synchronize()
        elif policy is not None:
            enqueue(False)
            torch.cuda.current_stream().synchronize()
        else:
            enqueue(True)
    for _ in range(warmup):
        one()
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    torch.cuda.synchronize(); barrier()
    el = time.perf_counter() - t0
    h2d = sum(v.numel() * 4 for k, v in hs_t.items() if k != "rigid_body_states") + act.numel() * 4
    h2d += N * 2 * 3 * 4                                  # rigid_body_states: the feet xyz columns only
    d2h = obs.numel() * 4 + (priv.numel() * 4 if P.num_priv_obs else 0) + rew.numel() * 4 + rst.numel() + tq.numel() * 4
    if policy is not None:
        d2h += act.numel() * 4
    return el, h2d, d2h


E2E_API = ("one CUDA-graph launch per step of: pf_step_host_begin + _join (C-ABI, pinned host buffers: PhysX state + actions in, "
           "obs/priv/rew/reset/torques out) + pf_rl_act on the device observations, sampled actions back to the host.  ONE bus "
           "round trip per POLICY step: the 4 torque sub-steps run on one uploaded dof_state, i.e. the leg assumes the physics "
           "step itself does not need the torques on the host between sub-steps (point_foot.py:180-185 is a host loop with "
           "gym.simulate inside; with PhysX on the host see e2e_substepped)")


def measure_workload(w, args, flush, barrier, peaks, device, min_seconds, e2e_steps, with_e2e=True):
    import torch
    import torch.distributed as dist
    ms_round, rounds, _ = time_rounds(w.step_fn, args.steps, args.warmup, flush, barrier, w.world, device, min_seconds)
    total = w.num_envs * w.world
    out = {"task": w.task, "num_envs": total, "num_envs_per_gpu": w.num_envs, "P": int(w.env._P.num_height_points),
           "ms_per_step": ms_round / args.steps, "value": total * args.steps / (ms_round * 1e-3), "rounds": rounds}
    post_r, act_r = rooflines(w, flush, peaks)
    out["roofline"] = {"post_physics": {k: post_r[k] for k in ("frac", "us_per_launch", "achieved", "traffic")},
                       "act": ({k: act_r[k] for k in ("bound", "hbm_frac", "tensor_frac", "us_per_launch")} if act_r else None)}
    ab = algorithmic_bytes_per_env(w.env._P)
    step_bytes = (ab["post_physics"] + ab["torques"] + (ab["act"] if w.policy is not None else 0)) * w.num_envs
    out["roofline"]["whole_step_hbm_frac"] = step_bytes / (out["ms_per_step"] * 1e-3) / 1e9 / peaks.get("hbm_gbs", 6650.0)
    if with_e2e:
        el, h2d, d2h = e2e_host(w, e2e_steps, 5, barrier, not args.no_graph)
        t = torch.tensor([el], device=device, dtype=torch.float64)
        if w.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["e2e"] = {"value": total * e2e_steps / float(t.item()), "unit": UNIT, "h2d_bytes_per_step": h2d,
                      "d2h_bytes_per_step": d2h, "steps": e2e_steps}
    return out, post_r, act_r


def runner_learn_fps(task, num_envs, device, iters=4, precision="bf16"):
    """OnPolicyRunner.learn() -- the loop a user of legged_gym runs -- timed by ITS OWN clocks (runner.log's Perf/collection time and
    Perf/learning_time, on_policy_runner.py:96-110): eager Python loop, one env.step + PPO.act + process_env_step per step, synthetic
    physics.  Median of ``iters`` iterations after one warm-up iteration."""
    from pointfoot_b200 import config, rl
    from pointfoot_b200.env import PointFoot
    cfg = config.PointFootFlatCfg() if task == "pointfoot_flat" else config.PointFootRoughCfg()
    cfg.env.num_envs = num_envs
    tcfg = config.class_to_dict(config.PointFootFlatCfgPPO() if task == "pointfoot_flat" else config.PointFootRoughCfgPPO())
    env = PointFoot(cfg, None, None, device, True, seed=1)
    old = os.environ.get("PF_RL_PRECISION")
    os.environ["PF_RL_PRECISION"] = precision
    try:
        runner = rl.OnPolicyRunner(env, tcfg, None, device)
    finally:
        if old is None:
            os.environ.pop("PF_RL_PRECISION", None)
        else:
            os.environ["PF_RL_PRECISION"] = old
    times = []
    for it in range(iters + 1):
        runner.learn(1)
        if it:
            times.append(runner.last_times)
    runner.alg.close()
    env.close()
    col = sorted(t[0] for t in times)[len(times) // 2]
    lrn = sorted(t[1] for t in times)[len(times) // 2]
    n = runner.num_steps_per_env * num_envs
    return {"task": task, "num_envs": num_envs, "steps_per_iteration": runner.num_steps_per_env, "precision": precision,
            "collection_s": col, "learning_s": lrn, "collection_env_steps_per_s": n / col, "total_fps": n / (col + lrn),
            "what": "pointfoot_b200.rl.OnPolicyRunner.learn(): wall clock of the runner's own timers, eager Python loop (not the CUDA-graph "
                    "step of `value`), synthetic physics"}


def ppo_iteration(n_per_gpu, task, precision, device, rank, world, iters=3):
    """BASELINE.json configs[3]: 24 steps x N envs of PPO.act + process_env_step on synthetic observations (the rollout's
    RL-side bookkeeping), compute_returns (GAE), update (5 epochs x 4 minibatches), CUDA events per phase, median of
    ``iters`` iterations after one warm-up iteration; MAX over ranks."""
    import torch
    import torch.distributed as dist
    from pointfoot_b200 import rl
    T = 24
    nc, hidden = (148, (512, 256, 128)) if task == "rough" else (27, (128, 64, 32))
    torch.manual_seed(0)
    ac = rl.ActorCritic(27, nc, 6, hidden, hidden, device=device)
    ppo = rl.PPO(ac, num_learning_epochs=5, num_mini_batches=4, gamma=0.99, lam=0.95, entropy_coef=0.01, schedule="adaptive",
                 desired_kl=0.01, precision=precision, device=device, env_id_offset=rank * n_per_gpu)
    ppo.init_storage(n_per_gpu, T, [27], [nc], [6])
    g = torch.Generator(device=device).manual_seed(rank)
    obs = torch.randn(T + 1, n_per_gpu, 27, device=device, generator=g)
    cobs = torch.randn(T + 1, n_per_gpu, nc, device=device, generator=g)
    rew = torch.randn(T, n_per_gpu, device=device, generator=g) * 0.1
    dones = torch.rand(T, n_per_gpu, device=device, generator=g) < 0.01
    tout = torch.zeros(n_per_gpu, dtype=torch.bool, device=device)
    res = []
    for it in range(iters + 1):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record()
        for t in range(T):
            ppo.act(obs[t], cobs[t])
            ppo.process_env_step(rew[t], dones[t], {"time_outs": tout})
        e[1].record()
        ppo.compute_returns(cobs[T])
        e[2].record()
        ppo.update()
        e[3].record()
        torch.cuda.synchronize()
        if it:
            res.append([e[i].elapsed_time(e[i + 1]) for i in range(3)])
    m = [sorted(r[i] for r in res)[len(res) // 2] for i in range(3)]
    out = {"launches_per_optimizer_step": ppo.launches_per_optimizer_step, "update_in_cuda_graph": ppo._update_graph is not None}
    if world > 1:
        mt = torch.tensor(m, device=device, dtype=torch.float64)
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        m = mt.tolist()
        # every rank must hold bit-identical parameters after the all-reduced steps
        ref = ac.flat.detach().clone()
        dist.broadcast(ref, 0)
        ok = torch.tensor([int(bool((ac.flat == ref).all().item()))], device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        out["parameters_identical_on_all_ranks"] = bool(ok.item())
        # the collectives alone: one gradient all-reduce (flat fp32 buffer) + one statistics all-reduce, as the update issues them
        n_ar = 50
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n_ar):
            dist.all_reduce(ppo.grads)
            dist.all_reduce(ppo.loss_stats)
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) / n_ar * 1e3], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["allreduce_us_per_optimizer_step"] = float(t.item())
        out["allreduce_bytes"] = int(ppo.grads.numel() * 4 + 64)
        out["allreduce_share_of_update"] = 20 * float(t.item()) * 1e-3 / m[2]
    fwd = mlp_flops_per_row(ac.actor_dims, ac.critic_dims) // 2
    upd_flops = 5 * T * n_per_gpu * world * 2 * fwd * 3          # fwd + dgrad + wgrad per sample per epoch
    out.update({"num_envs": n_per_gpu * world, "nets": task, "precision": precision, "rollout_policy_ms": m[0], "gae_ms": m[1],
                "update_ms": m[2], "iteration_ms": sum(m), "update_tflops": upd_flops / (m[2] * 1e-3) / 1e12})
    ppo.close()
    return out



def sweep_cases(world):
    """(task, envs per GPU, 187-point grid) of the size sweep.  One GPU: both tasks at 4,096 / 65,536 / 262,144 envs plus the
    rough task on the LeggedRobotCfg grid.  Several: C5 -- a fixed 262,144-env job sharded over the ranks (strong) and 32,768
    envs per GPU (weak); at 8 ranks the two are the same case and run once."""
    if world == 1:
        return [(t, n, g) for t, g in (("pointfoot_flat", False), ("pointfoot_rough", False), ("pointfoot_rough", True))
                for n in (4096, 65536, 262144)]
    return [(t, n, False) for t in ("pointfoot_flat", "pointfoot_rough") for n in dict.fromkeys((262144 // world, 32768))]

def run_b200(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    device = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(device))

    def barrier():
        if world > 1:
            dist.barrier()

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=device)

    def flush():
        if args.l2 == "flush":
            flush_buf.fill_(1)

    # ---- headline: configs[1] ----
    w = Workload(args.task, args.num_envs, device, rank, world, policy=not args.no_policy, graph=not args.no_graph)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    head, post_r, act_r = measure_workload(w, args, flush, barrier, peaks, device, args.min_seconds, min(max(args.steps, 50), 200))
    clocks = sampler.stop() if rank == 0 else None
    launches = w.launches_per_step
    policy_desc = w.policy.describe() if w.policy else "none"
    dtype = "f32" if w.policy is None else w.policy.dtype_name
    w.close()

    # the same workload with the fp32 policy (north_star: a stated looser tolerance where the MLP runs in bf16 -- 2e-2;
    # this is the line at fp32 precision beside it)
    value_fp32 = None
    if not args.no_policy:
        w32 = Workload(args.task, args.num_envs, device, rank, world, precision="fp32", graph=not args.no_graph)
        ms32, r32, _ = time_rounds(w32.step_fn, args.steps, args.warmup, flush, barrier, world, device, min(args.min_seconds, 0.25))
        value_fp32 = {"value": args.num_envs * world * args.steps / (ms32 * 1e-3), "ms_per_step": ms32 / args.steps,
                      "dtype": "f32", "policy": "fp32 SIMT GEMM chain (rtol 1e-4 vs torch fp32)"}
        w32.close()

    # ---- sweep (1 GPU) / config C5 (N GPUs) ----
    sweep = []
    if not args.no_sweep:
        cases = sweep_cases(world)
        terrain = None
        for task, n, g187 in cases:
            if world == 1 and task == args.task and n == args.num_envs and not g187:
                sweep.append(dict(head, scaling="headline"))
                continue
            if task == "pointfoot_rough" and terrain is None:
                terrain = rough_terrain(device)
            ws = Workload(task, n, device, rank, world, grid187=g187, terrain=terrain if task == "pointfoot_rough" else None)
            r, _, _ = measure_workload(ws, args, flush, barrier, peaks, device, min(args.min_seconds, 0.25), 30)
            if world > 1:
                r["scaling"] = ("strong (262,144 envs in total) = weak (32,768 envs per GPU)" if 262144 // world == 32768 else
                                "strong (262,144 envs in total)" if n == 262144 // world else "weak (32,768 envs per GPU)")
            sweep.append(r)
            ws.close()
            torch.cuda.empty_cache()

    # ---- PPO iteration: configs[3] ----
    ppo = []
    if not args.no_ppo:
        for nets in ("flat", "rough"):
            for prec in ("bf16", "fp32"):
                r = ppo_iteration(4096, nets, prec, device, rank, world)
                r["update_frac_of_bf16_sustained"] = r["update_tflops"] / peaks.get("bf16_tflops_sustained", 1381.4) if prec == "bf16" else None
                ppo.append(r)
        if world > 1:       # the update at scale: 32,768 envs per GPU, rough nets, bf16
            ppo.append(ppo_iteration(32768, "rough", "bf16", device, rank, world, iters=2))
        torch.cuda.empty_cache()

    # ---- the reference's own training loop: OnPolicyRunner.learn() with its own timers ----
    runner_learn = None
    if not args.no_ppo and world == 1:
        runner_learn = runner_learn_fps(args.task, args.num_envs, device)

    if rank == 0:
        total_envs = args.num_envs * world
        cfg = bench_config(args.task, args.num_envs, world, not args.no_policy)
        line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": dtype, "data": "synthetic", "config": cfg,
                "timing": {"rounds_of_K_steps": head["rounds"], "reported": "median round", "min_timed_seconds": args.min_seconds,
                           "cuda_graph": not args.no_graph, "policy_kernels": policy_desc,
                           "l2": "256 MB write before every timed step, excluded from the per-step CUDA-event time"
                                 if args.l2 == "flush" else "DIAGNOSTIC RUN, caches hot: not a valid bench line"},
                "clocks": clocks, "gpu_launches": launches * args.steps * head["rounds"], "gpu_launches_per_step": launches,
                "roofline": post_r, "roofline_act": act_r, "value_fp32": value_fp32,
                "e2e": dict(head["e2e"], api=E2E_API), "sweep": sweep, "ppo_iteration": ppo, "runner_learn": runner_learn}
        if not args.no_eager and world == 1:
            try:
                line["cuda_eager_baseline"] = cuda_eager_steps_per_sec(args.task, total_envs, device, flush)
                line["cuda_eager_baseline"]["speedup_of_value"] = head["value"] / line["cuda_eager_baseline"]["value"]
                r = cuda_eager_steps_per_sec("pointfoot_rough", 4096, device, flush, budget_s=3.0)
                line["cuda_eager_baseline"]["pointfoot_rough_4096"] = {k: r[k] for k in ("value", "ms_per_step", "steps")}
            except Exception as ex:           # reported, never silently dropped
                line["cuda_eager_baseline"] = {"unavailable": f"{type(ex).__name__}: {ex}"}
        if not args.no_cpu_baseline and world == 1:
            v, n, cores, kind = cpu_env_steps_per_sec(args.task, total_envs)
            what = "UNMODIFIED reference point_foot.py (baseline/_ref)" if kind == "reference" else "oracle port of the reference torch code"
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                                    "sample": f"{n} steps of {total_envs} envs ({args.task}), {what} + rsl_rl ActorCritic (restated), "
                                              "torch CPU eager, physics excluded"}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    if world > 1:
        from pointfoot_b200 import shard
        shard.shutdown()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
