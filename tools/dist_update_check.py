"""Multi-GPU PPO.update check (run under torchrun, one rank per GPU):
  1. the update captured into ONE CUDA graph together with its NCCL all-reduces (gradient AVG + loss statistics) gives
     bit-identical parameters to the same update launched step by step (PF_RL_UPDATE_GRAPH_DIST=0), on every rank, and all
     ranks end with the same parameters;
  2. a sharded OnPolicyRunner.learn() runs, and Episode/* statistics are merged over the shards.
usage: torchrun --nproc-per-node 2 tools/dist_update_check.py [num_envs_per_rank]"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import torch.distributed as dist

from pointfoot_b200 import config, rl
from pointfoot_b200.env import PointFoot

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
dev = f"cuda:{torch.cuda.current_device()}"
dist.init_process_group("nccl", device_id=torch.device(dev))
T, nc, hidden = 24, 148, (512, 256, 128)
out = {"world_size": world}


def run_update(graph_dist, precision):
    os.environ["PF_RL_UPDATE_GRAPH_DIST"] = "1" if graph_dist else "0"
    torch.manual_seed(0)
    ac = rl.ActorCritic(27, nc, 6, hidden, hidden, device=dev)
    ppo = rl.PPO(ac, num_learning_epochs=5, num_mini_batches=4, gamma=0.99, lam=0.95, entropy_coef=0.01, schedule="adaptive",
                 desired_kl=0.01, precision=precision, device=dev, env_id_offset=rank * N)
    ppo.init_storage(N, T, [27], [nc], [6])
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for t in range(T):
        ppo.act(torch.randn(N, 27, device=dev, generator=g), torch.randn(N, nc, device=dev, generator=g))
        ppo.process_env_step(torch.randn(N, device=dev, generator=g) * 0.1, torch.rand(N, device=dev, generator=g) < 0.02,
                             {"time_outs": torch.zeros(N, dtype=torch.bool, device=dev)})
    ppo.compute_returns(torch.randn(N, nc, device=dev, generator=g))
    losses = [ppo.update()]
    # a second update on the same storage (replay of the captured graph / a second eager pass)
    ppo.storage.step = T
    losses.append(ppo.update())
    torch.cuda.synchronize()
    flat = ac.flat.detach().clone()
    in_graph = ppo._update_graph is not None
    lr = ppo.learning_rate
    ppo.close()
    return flat, in_graph, losses, lr


for precision in ("fp32", "bf16"):
    pg, in_graph, lg, lrg = run_update(True, precision)
    pe, in_graph_e, le, lre = run_update(False, precision)
    ref = pg.clone()
    dist.broadcast(ref, 0)
    same_ranks = torch.tensor([int(torch.equal(pg, ref))], device=dev)
    dist.all_reduce(same_ranks, op=dist.ReduceOp.MIN)
    same_modes = torch.tensor([int(torch.equal(pg, pe))], device=dev)
    dist.all_reduce(same_modes, op=dist.ReduceOp.MIN)
    out[precision] = {"captured_update_in_graph": in_graph, "eager_update_in_graph": in_graph_e,
                      "parameters_identical_on_all_ranks": bool(same_ranks.item()),
                      "graph_equals_eager_bit_for_bit": bool(same_modes.item()),
                      "max_abs_diff_graph_vs_eager": float((pg - pe).abs().max()), "losses_graph": lg, "losses_eager": le,
                      "lr": (lrg, lre)}

# ---- sharded runner.learn(): Episode/* merged over the shards ----
os.environ["PF_RL_UPDATE_GRAPH_DIST"] = "1"
task = "pointfoot_rough"
cfg = config.PointFootRoughCfg()
cfg.env.num_envs = 1024
tcfg = config.class_to_dict(config.PointFootRoughCfgPPO())
env = PointFoot(cfg, None, None, dev, True, seed=1, env_id_offset=rank * 1024)
runner = rl.OnPolicyRunner(env, tcfg, None, dev)
seen = {}
orig = rl.shard.allreduce_episode_infos


def spy(ep_infos, num_envs):
    r = orig(ep_infos, num_envs)
    seen.update(r)
    return r


rl.shard.allreduce_episode_infos = spy
runner.learn(2, init_at_random_ep_len=True)
vals = torch.tensor([seen.get("rew_tracking_lin_vel", float("nan")), seen.get("terrain_level", float("nan"))], device=dev,
                    dtype=torch.float64)
lo, hi = vals.clone(), vals.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN)
dist.all_reduce(hi, op=dist.ReduceOp.MAX)
out["runner"] = {"episode_keys": len(seen), "merged_stats_identical_on_all_ranks": bool(torch.equal(lo, hi)),
                 "rew_tracking_lin_vel": float(vals[0]), "terrain_level": float(vals[1]),
                 "losses": [float(x) for x in runner.last_losses]}
if rank == 0:
    print(json.dumps(out), flush=True)
runner.alg.close()
env.close()
rl.shard.shutdown()
