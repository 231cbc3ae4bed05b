"""PPO iteration timing (BASELINE.json configs[3]): 24 steps x N envs of act + process_env_step on synthetic
observations, compute_returns (GAE), update (5 epochs x 4 minibatches), with CUDA events per phase.
usage: python tools/ppo_bench.py [num_envs] [task] [iters]      (PF_RL_PRECISION=fp32|bf16)"""
import json
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from pointfoot_b200 import rl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
task = sys.argv[2] if len(sys.argv) > 2 else "rough"
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
T = 24
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
if world > 1:                                  # torchrun: N envs per rank, NCCL all-reduce of the flat gradient per optimizer step
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
nc, hidden = (148, (512, 256, 128)) if task == "rough" else (27, (128, 64, 32))
precision = os.environ.get("PF_RL_PRECISION", "fp32")
torch.manual_seed(0)
dev = f"cuda:{torch.cuda.current_device()}"
ac = rl.ActorCritic(27, nc, 6, hidden, hidden, device=dev)
ppo = rl.PPO(ac, num_learning_epochs=5, num_mini_batches=4, gamma=0.99, lam=0.95, entropy_coef=0.01, schedule=os.environ.get("PF_SCHEDULE", "adaptive"),
             desired_kl=0.01, precision=precision, device=dev, env_id_offset=rank * N)
ppo.init_storage(N, T, [27], [nc], [6])
g = torch.Generator(device=dev).manual_seed(rank)
obs = torch.randn(T + 1, N, 27, device=dev, generator=g)
cobs = torch.randn(T + 1, N, nc, device=dev, generator=g)
rew = torch.randn(T, N, device=dev, generator=g) * 0.1
dones = torch.rand(T, N, device=dev, generator=g) < 0.01
tout = torch.zeros(N, dtype=torch.bool, device=dev)


def ev():
    return torch.cuda.Event(enable_timing=True)


res = []
for it in range(iters + 1):
    e = [ev() for _ in range(4)]
    e[0].record()
    for t in range(T):
        ppo.act(obs[t], cobs[t])
        ppo.process_env_step(rew[t], dones[t], {"time_outs": tout})
    e[1].record()
    ppo.compute_returns(cobs[T])
    e[2].record()
    vl, sl = ppo.update()
    e[3].record()
    torch.cuda.synchronize()
    if it:
        res.append([e[i].elapsed_time(e[i + 1]) for i in range(3)])
m = [sorted(r[i] for r in res)[len(res) // 2] for i in range(3)]
fwd = sum(a * b for a, b in zip(ac.actor_dims[:-1], ac.actor_dims[1:])) + sum(a * b for a, b in zip(ac.critic_dims[:-1], ac.critic_dims[1:]))
upd_flops = 5 * T * N * 2 * fwd * 3          # fwd + dgrad + wgrad per sample per epoch
if world > 1:
    flat = ac.flat.detach().clone()
    ref = flat.clone(); dist.broadcast(ref, 0)
    same = bool((flat == ref).all().item())      # every rank must hold bit-identical parameters after the all-reduced steps
    mt = torch.tensor(m, device=dev, dtype=torch.float64); dist.all_reduce(mt, op=dist.ReduceOp.MAX); m = mt.tolist()
    ok = torch.tensor([int(same)], device=dev); dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world_size": world, "parameters_identical_on_all_ranks": bool(ok.item())}))
    N *= world
if rank == 0:
  print(json.dumps({"num_envs": N, "task": task, "precision": precision, "rollout_policy_ms": m[0], "gae_ms": m[1], "update_ms": m[2],
                    "iteration_ms": sum(m), "update_tflops": upd_flops * world / m[2] / 1e9, "value_loss": vl, "surrogate_loss": sl, "learning_rate": ppo.learning_rate}))
if world > 1:
    dist.destroy_process_group()
