"""Times the LeggedRobot (a1) step path on one GPU: pf_lr_pd_torques and pf_lr_post_physics_step (scan + step + height columns
+ finalize), CUDA events, L2 flushed between calls, against the HBM bound of the algorithmic bytes.
usage: python tools/lr_bench.py [num_envs ...]"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

from pointfoot_b200 import config, synth
from pointfoot_b200.legged_robot import LeggedRobot

dev = "cuda:0"
sizes = [int(a) for a in sys.argv[1:]] or [4096, 65536, 262144]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
try:
    hbm = float(json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except (OSError, ValueError, KeyError):
    hbm = 6650.0


def timed(fn, n=10):
    tot = 0.0
    for i in range(n + 3):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        if i >= 3:
            tot += a.elapsed_time(b)
    return tot / n


terrain = synth.make_heightfield(1)
out = {"hbm_peak_gbs": hbm, "rows": []}
for n in sizes:
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = n
    env = LeggedRobot(cfg, None, None, dev, True, state=synth.make_lr_state(n, seed=1, terrain=terrain))
    P = env._P
    D, R, Pn = P.num_dof, P.num_reward_terms + P.has_termination_term, P.num_height_points
    # per env: PhysX rows in (root 52, dof 8D, contacts 12B), env-owned state in and out, observations out
    read = 52 + 8 * D + 12 * P.num_bodies + 16 + 8 + 4 * D * 3 + 4 * P.num_feet + P.num_feet + 4 * R
    write = 4 * P.num_obs + 36 + 4 * D * 2 + 24 + 4 * P.num_feet + P.num_feet + 16 + 8 + 4 + 2 + 4 * R
    step_bytes = (read + write) * n
    a = torch.randn(n, 12, device=dev)
    env.step(a)
    t_post = timed(lambda: env._call_post_physics_step(0))
    t_tq = timed(lambda: env._call_pd_torques(a))
    out["rows"].append({"num_envs": n, "post_physics_us": t_post * 1e3, "algorithmic_bytes": step_bytes,
                        "post_physics_GBs": step_bytes / t_post / 1e6, "post_physics_frac_of_hbm": step_bytes / t_post / 1e6 / hbm,
                        "pd_torques_us": t_tq * 1e3, "pd_torques_frac_of_hbm": (16 * D * n) / t_tq / 1e6 / hbm})
    env.close()
print(json.dumps(out))
