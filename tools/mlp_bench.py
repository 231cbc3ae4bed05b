"""Times the fused tcgen05 MLP forward alone (actor + critic + sample = PPO.act) with CUDA events.
usage: python tools/mlp_bench.py [num_envs] [task] [iters]      (MLP_NO / MLP_NC: actor / critic input widths)"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from pointfoot_b200 import rl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
task = sys.argv[2] if len(sys.argv) > 2 else "rough"
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
nc, hidden = (148, (512, 256, 128)) if task == "rough" else (27, (128, 64, 32))
no, nc = int(os.environ.get("MLP_NO", 27)), int(os.environ.get("MLP_NC", nc))
ac = rl.ActorCritic(no, nc, 6, hidden, hidden)
ppo = rl.PPO(ac, precision=os.environ.get("PF_RL_PRECISION", "bf16"), num_mini_batches=4, num_learning_epochs=1)
ppo.init_storage(N, 2, [no], [nc], [6])
obs, cobs = torch.randn(N, no, device="cuda"), torch.randn(N, nc, device="cuda")
flops = 2 * N * (sum(a * b for a, b in zip(ac.actor_dims[:-1], ac.actor_dims[1:])) + sum(a * b for a, b in zip(ac.critic_dims[:-1], ac.critic_dims[1:])))
for _ in range(3):
    ppo.storage.step = 0; ppo.act(obs, cobs)
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
for a, b in ev:
    ppo.storage.step = 0
    a.record(); ppo.act(obs, cobs); b.record()
torch.cuda.synchronize()
ms = sorted(a.elapsed_time(b) for a, b in ev)[iters // 2]
print(f"N={N} {task}: PPO.act {ms*1e3:.1f} us  {flops/ms/1e9:.1f} TFLOP/s (algorithmic, {flops/1e9:.1f} GFLOP)")
