"""GPU debug of the tcgen05 fused MLP: max abs error of actor / critic outputs against torch fp32 on bf16-rounded
operands, for a few shapes.  usage: [PF_TC_SWAP=1] python tools/tc_debug.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from pointfoot_b200 import rl, capi

def ref_mlp(ac, name, x):
    dims = ac.actor_dims if name == "actor" else ac.critic_dims
    h = x
    for l in range(len(dims) - 1):
        W, b = ac._views[f"{name}.{2*l}.weight"], ac._views[f"{name}.{2*l}.bias"]
        h = h.to(torch.bfloat16).float() @ W.to(torch.bfloat16).float().t() + b
        if l < len(dims) - 2:
            h = torch.nn.functional.elu(h)
    return h

for (N, nc, hidden) in [(128, 27, (128, 64, 32)), (300, 148, (512, 256, 128)), (4096, 148, (512, 256, 128))]:
    ac = rl.ActorCritic(27, nc, 6, hidden, hidden)
    ppo = rl.PPO(ac, precision="bf16", num_mini_batches=4, num_learning_epochs=1)
    ppo.init_storage(N, 4, [27], [nc], [6])
    torch.manual_seed(0)
    obs, cobs = torch.randn(N, 27, device="cuda"), torch.randn(N, nc, device="cuda")
    mean = ac.act_inference(obs); val = ac.evaluate(cobs); torch.cuda.synchronize()
    rm, rv = ref_mlp(ac, "actor", obs), ref_mlp(ac, "critic", cobs)
    print(f"N={N} hidden={hidden}: actor max|err|={float((mean-rm).abs().max()):.3e} (|ref| {float(rm.abs().max()):.3f})  "
          f"critic max|err|={float((val-rv).abs().max()):.3e} (|ref| {float(rv.abs().max()):.3f})", flush=True)
    ppo.close()
