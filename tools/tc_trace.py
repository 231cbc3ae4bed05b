"""Role timelines of CTA 0 of the fused tcgen05 MLP kernel (pf_rl_tc_trace): where one work item spends its time.
usage: python tools/tc_trace.py [num_envs] [task] [max_events]"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from pointfoot_b200 import rl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
task = sys.argv[2] if len(sys.argv) > 2 else "rough"
limit = int(sys.argv[3]) if len(sys.argv) > 3 else 140
nc, hidden = (148, (512, 256, 128)) if task == "rough" else (27, (128, 64, 32))
ac = rl.ActorCritic(27, nc, 6, hidden, hidden)
ppo = rl.PPO(ac, precision="bf16", num_mini_batches=4, num_learning_epochs=1)
ppo.init_storage(N, 2, [27], [nc], [6])
obs, cobs = torch.randn(N, 27, device="cuda"), torch.randn(N, nc, device="cuda")
for _ in range(3):
    ppo.storage.step = 0; ppo.act(obs, cobs)
CAP = 512
buf = torch.zeros(4 * CAP, dtype=torch.int64, device="cuda")
assert ppo._lib.pf_rl_tc_trace(ppo._ctx, C.c_void_p(buf.data_ptr())) == 0
ppo.storage.step = 0; ppo.act(obs, cobs)
torch.cuda.synchronize()
ppo._lib.pf_rl_tc_trace(ppo._ctx, None)
raw = buf.cpu().tolist()
NAMES = {0x1: "load  group", 0x2: "mma   reach", 0x3: "mma   issue", 0x4: "mma   done ", 0x5: "epi   item ", 0x6: "epi   reach",
         0x7: "epi   start", 0x8: "epi   work ", 0x9: "epi   done ", 0xA: "mma   tile wait", 0xB: "mma   tile got", 0xC: "epi   tmem ld", 0xD: "epi   tmem got"}
ev = []
for role in range(4):
    for v in raw[role * CAP:(role + 1) * CAP]:
        if v:
            tag, clk = (v >> 48) & 0xFFFF, v & ((1 << 48) - 1)
            ev.append((clk, role, tag))
ev.sort()
t0 = ev[0][0]
prev = {}
for clk, role, tag in ev[:limit]:
    d = clk - prev.get(role, t0)
    prev[role] = clk
    sub = {0x500: "begin", 0x501: "staged", 0x502: "arrived", 0x503: "loaded", 0x504: "padded", 0x505: "signalled"}.get(tag, str(tag & 0xFF))
    print(f"{(clk - t0):9d} cyc  (+{d:7d})  {'  ' * role}{NAMES[tag >> 8]} {sub}")
last = max(c for c, r, t in ev if r == 0)
items = sum(1 for c, r, t in ev if t == 0x500)
print(f"CTA 0: {items} items traced, {(last - t0) / max(items, 1):.0f} cycles per item")
