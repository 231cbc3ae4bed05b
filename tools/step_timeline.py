"""Device timeline of one CUDA-graphed policy step (pf_debug_timeline / pf_rl_debug_timeline): when each kernel's CTA 0
starts, passes its griddepcontrol.wait and ends, in ns of the GPU's global timer.  Needs the diagnostic build:
    PF_BUILD_TIMELINE=1 python __graft_entry__.py --force      (the marks are compiled out of the product build)
usage: python tools/step_timeline.py [task] [num_envs]"""
import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                        # noqa: E402
import bench                                        # noqa: E402
from pointfoot_b200.env import GraphedStep          # noqa: E402

task = sys.argv[1] if len(sys.argv) > 1 else "pointfoot_flat"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
device = torch.device("cuda:0")
env, cfg, train_cfg, state = bench.build_env(task, n, device, 0, 1)
args = argparse.Namespace(no_policy=False, precision=None)
policy = bench.make_policy(env, train_cfg, args)
actions = policy.actions
env.reset()
step = GraphedStep(env, actions, lambda: policy.act_into(env, actions))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
log = torch.zeros(4096, dtype=torch.int64, device=device)
lib = env._lib
for _ in range(5):
    flush.fill_(1); step()
torch.cuda.synchronize()
if lib.pf_debug_timeline(C.c_void_p(log.data_ptr())) != 0 or lib.pf_rl_debug_timeline(C.c_void_p(log.data_ptr())) != 0:
    sys.exit("timeline marks are not compiled in: PF_BUILD_TIMELINE=1 python __graft_entry__.py --force")
for _ in range(3):
    log[4000:4002] = 0                 # per-step maxima (the last step's survive)
    flush.fill_(1); step()
torch.cuda.synchronize()
lib.pf_debug_timeline(None); lib.pf_rl_debug_timeline(None)
raw = log.cpu().tolist()
ev = sorted(((v & ((1 << 48) - 1)), (v >> 48) & 0xFFFF) for v in raw[1:1 + raw[0]])
NAMES = {0x10: "torques start", 0x11: "torques past wait", 0x12: "torques end", 0x20: "post start", 0x21: "post past wait", 0x22: "post end (CTA 0)",
         0x30: "finalize start", 0x31: "finalize past wait", 0x32: "finalize ids written", 0x40: "scan start", 0x41: "scan past wait", 0x42: "scan end (CTA 0)",
         0x50: "act start", 0x51: "act prologue done", 0x52: "act past wait", 0x53: "act end (CTA 0)"}
# last step = events after the last big gap (the L2 flush)
cut = 0
for i in range(1, len(ev)):
    if ev[i][0] - ev[i - 1][0] > 20000:
        cut = i
t0 = ev[cut][0]
for t, tag in ev[cut:]:
    TC = {0x1: "loader: group", 0x2: "mma: reached group", 0x3: "mma: issue group", 0x4: "mma: group done", 0x5: "stage", 0x6: "epilogue: reached", 0x7: "epilogue: accumulator ready",
          0x8: "epilogue: written", 0x9: "epilogue: done", 0xA: "mma: tile wait", 0xB: "mma: tile got", 0xC: "epi: tmem ld", 0xD: "epi: tmem got"}
    name = NAMES.get(tag)
    if name is None and tag >= 0x1000:
        k, sub = (tag - 0x1000) >> 8, (tag - 0x1000) & 0xFF
        if k in (0xC, 0xD):
            continue
        name = "   act " + TC.get(k, hex(k)) + " " + ({0: "begin", 2: "arrived", 3: "loaded", 4: "padded", 5: "signalled"}.get(sub, str(sub)) if k == 5 else str(sub))
    name = name or hex(tag)
    print(f"{(t - t0) / 1000:8.2f} us  {name}")
if raw[4000] or raw[4001]:
    print(f"step kernel, last warp to finish: without a reset env {(raw[4000] - t0) / 1000:8.2f} us, with one {(raw[4001] - t0) / 1000:8.2f} us")
