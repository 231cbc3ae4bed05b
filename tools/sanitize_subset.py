"""The subset of the GPU path run under compute-sanitizer (memcheck / racecheck / synccheck; profiles/r2_sanitizer_*.txt):
every kernel family once at a given env count -- env step (flat + rough, with resets and a push), PPO.act (fp32 and the
tcgen05 kernel), GAE, one PPO.update (fp32 and tcgen05), the LeggedRobot step, the actuator network, terrain generation.
usage: python tools/sanitize_subset.py <num_envs> [parts]   (parts: comma list of env,rl,lr,act,terrain; default all)"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import torch

N = int(sys.argv[1]) if len(sys.argv) > 1 else 33
parts = set((sys.argv[2] if len(sys.argv) > 2 else "env,rl,lr,act,terrain").split(","))
dev = "cuda:0"
from pointfoot_b200 import config, rl, synth                     # noqa: E402

if "env" in parts:
    import common
    from pointfoot_b200.env import PointFoot
    for task in ("pointfoot_flat", "pointfoot_rough"):
        state = synth.make_state(N, task, seed=3, adversarial=True)
        cfg = common.make_cfg(task, N)
        cfg.domain_rand.push_interval_s = 0.04               # a push on the 2nd step
        env = PointFoot(cfg, None, None, dev, True, state=common.clone_state(state), seed=1, keep_measured_heights=(task == "pointfoot_rough"))
        a = torch.randn(N, 6, device=dev)
        env.reset()
        for _ in range(3):
            env.step(a)
        torch.cuda.synchronize()
        env.close()
        print("env", task, N, "ok", flush=True)

if "rl" in parts:
    T = 4
    os.environ["PF_RL_UPDATE_GRAPH"] = "0"                    # eager launches: the sanitizer sees every kernel
    for precision, hidden in (("fp32", (64, 32)), ("bf16", (512, 256, 128)), ("bf16", (128, 64, 32))):
        torch.manual_seed(0)
        ac = rl.ActorCritic(27, 148, 6, hidden, hidden, device=dev)
        ppo = rl.PPO(ac, num_learning_epochs=1, num_mini_batches=2, precision=precision, device=dev)
        ppo.init_storage(N, T, [27], [148], [6])
        for t in range(T):
            ppo.act(torch.randn(N, 27, device=dev), torch.randn(N, 148, device=dev))
            ppo.process_env_step(torch.randn(N, device=dev) * 0.1, torch.rand(N, device=dev) < 0.1,
                                 {"time_outs": torch.zeros(N, dtype=torch.bool, device=dev)})
        ppo.compute_returns(torch.randn(N, 148, device=dev))
        ppo.update()
        ac.act_inference(torch.randn(N, 27, device=dev))
        torch.cuda.synchronize()
        ppo.close()
        print("rl", precision, hidden, N, "ok", flush=True)

if "lr" in parts:
    from pointfoot_b200.legged_robot import LeggedRobot
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = N
    cfg.domain_rand.push_interval_s = 0.04
    env = LeggedRobot(cfg, None, None, dev, True, state=synth.make_lr_state(N, seed=3, adversarial=True))
    env.reset()
    for _ in range(3):
        env.step(torch.randn(N, 12, device=dev))
    torch.cuda.synchronize()
    print("lr", N, "ok", flush=True)

if "act" in parts:
    from pointfoot_b200.anymal import ActuatorNetwork
    from pointfoot_b200 import capi
    net = ActuatorNetwork(N, 12, 0.5, torch.zeros(1, 12, device=dev), dev)
    net.weights = (torch.randn(capi.ACTUATOR_NET_WEIGHTS, device=dev) * 0.3).contiguous()
    dof = torch.randn(N * 12, 2, device=dev)
    for _ in range(2):
        net.compute_torques(torch.randn(N, 12, device=dev), dof)
    ids = torch.arange(min(N, 5), dtype=torch.int32, device=dev)
    pad = torch.zeros(N, dtype=torch.int32, device=dev); pad[:ids.numel()] = ids
    net.reset(pad, torch.tensor([ids.numel()], dtype=torch.int32, device=dev))
    torch.cuda.synchronize()
    print("actuator", N, "ok", flush=True)

if "terrain" in parts:
    import parity_terrain
    from pointfoot_b200.terrain import Terrain
    t = Terrain(parity_terrain.make_cfg(parity_terrain.t_all_kinds_random), N, device=dev, seed=1)
    torch.cuda.synchronize()
    print("terrain", tuple(t.height_field_raw.shape), "ok", flush=True)
print("subset ok", N)
