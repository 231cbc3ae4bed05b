"""Role timeline of CTA 0 of pf_train_gemm_kernel over one minibatch step (pf_tt_debug_timeline).  Needs the diagnostic
build:  PF_BUILD_TIMELINE=1 python __graft_entry__.py --force   (the marks are compiled out of the product build)
usage: python tools/train_trace.py [num_envs] [rough|flat]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from pointfoot_b200 import rl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
task = sys.argv[2] if len(sys.argv) > 2 else "rough"
nc, hidden = (148, (512, 256, 128)) if task == "rough" else (27, (128, 64, 32))
os.environ["PF_RL_UPDATE_GRAPH"] = "0"
T = 24
ac = rl.ActorCritic(27, nc, 6, hidden, hidden)
ppo = rl.PPO(ac, precision="bf16", num_mini_batches=4, num_learning_epochs=1)
ppo.init_storage(N, T, [27], [nc], [6])
for t in range(T):
    ppo.act(torch.randn(N, 27, device="cuda"), torch.randn(N, nc, device="cuda"))
    ppo.process_env_step(torch.randn(N, device="cuda") * 0.1, torch.rand(N, device="cuda") < 0.02,
                         {"time_outs": torch.zeros(N, dtype=torch.bool, device="cuda")})
ppo.compute_returns(torch.randn(N, nc, device="cuda"))
ppo.update()                                   # warm-up
ppo.storage.step = T
log = torch.zeros(4096, dtype=torch.int64, device="cuda")
lib = ppo._lib
if lib.pf_tt_debug_timeline(C.c_void_p(log.data_ptr())) != 0:
    sys.exit("timeline marks are not compiled in: PF_BUILD_TIMELINE=1 python __graft_entry__.py --force")
ppo.update()
torch.cuda.synchronize()
lib.pf_tt_debug_timeline(None)
raw = log.cpu().tolist()
ev = sorted(((v & ((1 << 48) - 1)), (v >> 48) & 0xFFFF) for v in raw[1:1 + min(raw[0], 4094)])
NAMES = {0x300: "load  tile (fwd)", 0x301: "load  tile (dgrad)", 0x302: "load  tile (wgrad)", 0x310: "mma   tile begin", 0x311: "mma   acc free",
         0x312: "mma   first stage landed", 0x313: "mma   last stage landed", 0x320: "epi   wait acc", 0x321: "epi   acc complete",
         0x330: "epi   done", 0x331: "mma   done", 0x332: "load  done", 0x340: "cta   prologue done", 0x341: "cta   past dependency wait"}
t0 = ev[0][0] if ev else 0
last = t0
print(f"{raw[0]} records; first minibatch step of the update (8 GEMM launches = 8 blocks between 'prologue done' marks)")
k = 0
for t, tag in ev:
    if tag == 0x340:
        k += 1
        if k > 9:
            break
        print(f"--- launch {k}")
    print(f"{(t - t0) / 1e3:10.2f} us  (+{(t - last) / 1e3:7.2f})  {NAMES.get(tag, hex(tag))}")
    last = t
ppo.close()
