"""Headline benchmark: env-steps/s of the fused step (4x PD torques + post_physics_step [+ ActorCritic act]) on
synthetic state tensors.  Contract: ``python bench.py --gpus N --steps K --warmup W`` prints ONE JSON line
(rank 0).  ``--impl reference`` times the reference's CPU implementation of the same path (the oracle port:
the reference is pure Python over a closed physics binary and cannot travel to the GPU box).

A "step" is one pass of the hot path over one batch of N envs per GPU: clip actions, decimation x
_compute_torques, post_physics_step (point_foot.py:170-230) and, when the RL kernels are built, the ActorCritic
act/evaluate that produces the next actions.  Inputs are resident in HBM for ``value``; ``e2e`` goes through the
host-buffer C-ABI call (pf_step_host) with the PhysX state tensors, actions and results crossing PCIe every step.
"""
import argparse
import ctypes as C
import json
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
    p.add_argument("--l2", default="flush", choices=["flush", "none"],
                   help="'none' is a DIAGNOSTIC (hot caches): such a line is marked invalid")
    return p.parse_args()


def make_cfgs(task, num_envs):
    from pointfoot_b200 import config
    cfg = getattr(config, TASKS[task][0])()
    train_cfg = getattr(config, TASKS[task][1])()
    cfg.env.num_envs = num_envs
    return cfg, train_cfg


def algorithmic_bytes_per_env(P):
    """compulsory fp32 traffic of one policy step, per env.
    post_physics_step: READ root 52 + dof_state 48 + contact 108 + feet xyz 24 + actions 24 + torques 24 +
    last_actions 24 + last_dof_vel 24 + commands 4C + ep_len 8 + feet state machine 38 + episode_sums 4R;
    WRITE obs 108 + priv 4*PO + rew 4 + flags 3 + base vel/gravity 36 + last_* 72 + ep_len 8 + feet SM 46 +
    episode_sums 4R.  Torques: 4 x (actions 24 + dof_state 48 + torques 24)."""
    C_, R = P.num_commands, P.num_reward_terms + P.has_termination_term
    read = 52 + 48 + 108 + 24 + 24 + 24 + 24 + 24 + 4 * C_ + 8 + 38 + 4 * R
    write = 4 * P.num_obs + 4 * P.num_priv_obs + 4 + 3 + 36 + 72 + 8 + 46 + 4 * R
    return {"post_physics": read + write, "torques": P.decimation * 96}


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


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_env_steps_per_sec(task, num_envs, budget_s=12.0, min_steps=5, threads=None, with_policy=True):
    """The oracle port of the reference path (oracle/pointfoot_ref.py, plain torch CPU ops in the reference's op
    order + oracle/rsl_rl_ref.py ActorCritic) timed on the host cores.  Returns (env-steps/s, steps, cores)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common
    from pointfoot_b200 import synth
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    state = synth.make_state(num_envs, task, seed=1)
    orc, (cfg, P, noise, pts, names) = common.oracle_env(task, state)
    policy = None
    if with_policy:
        try:
            from oracle import rsl_rl_ref
            _, train_cfg = make_cfgs(task, num_envs)
            torch.manual_seed(1)
            policy = rsl_rl_ref.ActorCritic(P.num_obs, P.num_priv_obs or P.num_obs, P.num_dof,
                                            train_cfg.policy.actor_hidden_dims, train_cfg.policy.critic_hidden_dims)
        except ImportError:
            policy = None
    actions = state["actions"].clone()

    def one():
        nonlocal actions
        obs, priv, rew, reset = orc.step(actions)
        if policy is not None:
            with torch.inference_mode():
                actions = policy.act(obs).detach()
                policy.evaluate(priv if priv is not None else obs)
                policy.get_actions_log_prob(actions)
    for _ in range(3):
        one()
    n, t0 = 0, time.perf_counter()
    while True:
        one()
        n += 1
        el = time.perf_counter() - t0
        if (el > budget_s and n >= min_steps) or n >= 100000:
            break
    return num_envs * n / el, n, cores, policy is not None


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path, all host threads, same config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps)
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common
    from pointfoot_b200 import synth
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_total = args.num_envs * args.gpus
    state = synth.make_state(n_total, args.task, seed=1)
    orc, (cfg, P, noise, pts, names) = common.oracle_env(args.task, state)
    policy = None
    try:
        from oracle import rsl_rl_ref
        _, train_cfg = make_cfgs(args.task, n_total)
        torch.manual_seed(1)
        policy = rsl_rl_ref.ActorCritic(P.num_obs, P.num_priv_obs or P.num_obs, P.num_dof,
                                        train_cfg.policy.actor_hidden_dims, train_cfg.policy.critic_hidden_dims)
    except ImportError:
        pass
    actions = state["actions"].clone()

    def one():
        nonlocal actions
        obs, priv, rew, reset = orc.step(actions)
        if policy is not None:
            with torch.inference_mode():
                actions = policy.act(obs).detach()
                policy.evaluate(priv if priv is not None else obs)
                policy.get_actions_log_prob(actions)
    for _ in range(max(3, args.warmup)):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    el = time.perf_counter() - t0
    v = n_total * steps / el
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": max(3, args.warmup), "ms_per_step": 1e3 * el / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.task} {n_total} envs: fused PD torques + post_physics_step + reward/obs"
                                   + (" + ActorCritic act/evaluate" if policy is not None else ""),
                       "num_envs": n_total, "task": args.task, "physics": "excluded (Isaac Gym not installable)"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{steps} steps of {n_total} envs, torch CPU eager, {cores} threads"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
def build_env(task, num_envs, device, rank, world, terrain=None):
    import torch
    from pointfoot_b200 import synth
    from pointfoot_b200.env import PointFoot
    cfg, train_cfg = make_cfgs(task, num_envs)
    state = synth.make_state(num_envs, task, seed=1, env_id_offset=rank * num_envs, terrain=terrain)
    if task == "pointfoot_rough":     # terrain_types is a function of the GLOBAL env id (point_foot.py:981-983)
        import numpy as np
        gid = np.arange(num_envs) + rank * num_envs
        state["terrain_types"] = torch.from_numpy(np.floor(gid / (num_envs * world / 20)).astype(np.int64))
    env = PointFoot(cfg, None, None, device, True, state=state, seed=train_cfg.seed, env_id_offset=rank * num_envs)
    return env, cfg, train_cfg, state


def make_policy(env, train_cfg, args):
    if args.no_policy:
        return None
    try:
        from pointfoot_b200 import rl
    except ImportError:
        return None
    return rl.make_bench_policy(env, train_cfg)


def time_steps(step_fn, steps, warmup, flush, barrier):
    """W untimed warm-up steps, then K timed steps.  L2 is flushed (256 MB write) before every timed step and the
    flush is excluded: each step is bracketed by its own pair of CUDA events on the launching stream."""
    import torch
    for _ in range(warmup):
        flush()
        step_fn()
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
    return sum(a.elapsed_time(b) for a, b in evs)          # ms over the K steps


def kernel_roofline(env, steps, flush, peaks, task):
    """Device time of the dominant kernel (pf_post_physics_kernel + its finalize) alone, CUDA events, L2 flushed."""
    import torch
    from pointfoot_b200 import capi
    lib, ctx, b = env._lib, env._ctx, C.byref(env._bufs)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    tot = 0.0
    n = max(10, min(steps, 50))
    for i in range(n + 3):
        flush()
        a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        capi.check(lib.pf_post_physics_step(ctx, b, C.c_int64(0), st), "pf_post_physics_step")
        e.record()
        torch.cuda.synchronize()
        env.common_step_counter += 1
        if i >= 3:
            tot += a.elapsed_time(e)
    ms = tot / n
    nbytes = algorithmic_bytes_per_env(env._P)["post_physics"] * env.num_envs
    achieved = nbytes / (ms * 1e-3) / 1e9
    peak = peaks.get("hbm_gbs", 6650.0)
    traffic = None
    try:        # measured DRAM bytes of the same kernel on the same workload, from the committed ncu capture
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f).get(f"{task}:{env.num_envs}", {}).get("bytes")
    except (OSError, ValueError):
        pass
    return {"bound": "hbm", "kernel": "pf_post_physics_kernel(+pf_finalize_kernel" + (", pf_height_scan_kernel" if env._P.num_height_points > 0 else "") + ")", "achieved": achieved,
            "peak": peak, "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback",
            "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "us_per_launch": ms * 1e3,
            "algorithmic_bytes_per_launch": nbytes}


def e2e_host(env, steps, warmup, barrier, policy=None, use_graph=True):
    """Same metric through the host-buffer C-ABI call: PhysX state tensors + actions H2D, results D2H, per step."""
    import torch
    from pointfoot_b200 import capi
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
        # the whole host-facing step (H2D copies from the pinned buffers, kernels, D2H copies) as ONE graph launch:
        # what a caller stepping a fixed set of pinned PhysX buffers does to get the ~15 API calls off the CPU path
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
    h2d += hs_t["rigid_body_states"].numel() * 4 if os.environ.get("PF_HOST_RBS", "") == "full" else N * 2 * 3 * 4   # feet xyz only
    d2h = obs.numel() * 4 + (priv.numel() * 4 if P.num_priv_obs else 0) + rew.numel() * 4 + rst.numel() + tq.numel() * 4
    if policy is not None:
        d2h += act.numel() * 4
    return el, h2d, d2h


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

    env, cfg, train_cfg, state = build_env(args.task, args.num_envs, device, rank, world)
    policy = make_policy(env, train_cfg, args)
    actions = policy.actions if policy is not None else torch.zeros(env.num_envs, env.num_actions, device=device)
    extra = (lambda: policy.act_into(env, actions)) if policy is not None else None
    # decimation x pf_pd_torques, [pf_height_scan_kernel when the task measures heights], pf_post_physics_kernel, pf_finalize_kernel
    launches_per_step = (cfg.control.decimation + 2 + (1 if env._P.num_height_points > 0 else 0)
                         + (policy.launches_per_act if policy is not None else 0))
    env.reset()
    if args.no_graph:
        def step_fn():
            env.step(actions)
            if extra:
                extra()
    else:
        from pointfoot_b200.env import GraphedStep
        step_fn = GraphedStep(env, actions, extra)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = time_steps(step_fn, args.steps, args.warmup, flush, barrier)
    t = torch.tensor([ms], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clocks = sampler.stop() if rank == 0 else None

    roof = kernel_roofline(env, args.steps, flush, peaks, args.task)
    e2e_s, h2d, d2h = e2e_host(env, min(args.steps, 200), max(3, min(args.warmup, 10)), barrier, policy, not args.no_graph)
    t = torch.tensor([e2e_s], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_steps = min(args.steps, 200)
    e2e_val = env.num_envs * world * e2e_steps / float(t.item())

    if rank == 0:
        total_envs = env.num_envs * world
        line = {"metric": METRIC, "value": total_envs * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32" if policy is None else policy.dtype_name,
                "data": "synthetic",
                "config": {"workload": f"{args.task} {env.num_envs} envs/GPU: fused PD torques x{cfg.control.decimation} + "
                                       "post_physics_step + reward/obs" + (" + ActorCritic act/evaluate" if policy else ""),
                           "task": args.task, "num_envs_per_gpu": env.num_envs, "num_envs": total_envs,
                           "policy": policy.describe() if policy else "none",
                           "physics": "excluded (Isaac Gym not installable): synthetic state tensors resident in HBM",
                           "cuda_graph": not args.no_graph,
                           "l2": "256 MB write before every timed step, excluded from the per-step CUDA-event time"
                                 if args.l2 == "flush" else "DIAGNOSTIC RUN, caches hot: not a valid bench line",
                           "parallelism": f"env-sharded x{world}, no data-path collective"},
                "clocks": clocks, "gpu_launches": launches_per_step * args.steps, "roofline": roof,
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "api": ("one CUDA-graph launch per step of: " if (policy and not args.no_graph) else "") + "pf_step_host / pf_step_host_begin + _join (C-ABI, pinned host buffers: PhysX state + actions in, obs/priv/rew/reset/torques out)" + (" + pf_rl_act on the device observations, sampled actions back to the host" if policy else ""),
                        "steps": e2e_steps}}
        if not args.no_cpu_baseline and world == 1:
            v, n, cores, with_pol = cpu_env_steps_per_sec(args.task, env.num_envs)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{n} steps of {env.num_envs} envs ({args.task}), oracle port of the "
                                              f"reference torch code on CPU{' + ActorCritic' if with_pol else ''}, physics excluded"}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    env.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
