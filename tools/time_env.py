"""Quick device timing of the env kernels (development tool; bench.py is the contract)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity                                  # noqa: E402
from pointfoot_b200 import synth, capi         # noqa: E402


def main():
    sizes = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["4096", "65536", "262144"])]
    tasks = sys.argv[2].split(",") if len(sys.argv) > 2 else ["pointfoot_flat", "pointfoot_rough"]
    for task in tasks:
        terrain = synth.make_heightfield(1) if task == "pointfoot_rough" else None
        for n in sizes:
            st = synth.make_state(n, task, seed=1, terrain=terrain)
            env = parity.EnvUnderTest(task, st, "cuda")
            if os.environ.get("PF_KEEP_MH", "0") == "0":       # as pointfoot_b200.env.PointFoot does by default: measured_heights not materialised
                env.t["measured_heights"] = None
                env.b.measured_heights = None
            a = st["actions"].cuda()
            lib, ctx, b = env.lib, env.ctx, env.b
            stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            def torq():
                lib.pf_pd_torques(ctx, C.byref(b), C.c_void_p(a.data_ptr()), stream)
            k = [0]
            def post():
                k[0] += 1
                lib.pf_post_physics_step(ctx, C.byref(b), C.c_int64(k[0]), stream)
            for name, fn in (("pd_torques", torq), ("post_physics+finalize", post)):
                for _ in range(5):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                iters = 50
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record(); torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / iters
                print(f"{task} N={n} {name}: {us:.2f} us  resets={int(env.t['reset_count'])}", flush=True)
            env.close()


if __name__ == "__main__":
    main()
