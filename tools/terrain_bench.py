"""Times terrain generation on one GPU: pf_terrain_generate (raster + env origins) and pf_terrain_trimesh on the default
LeggedRobotCfg.terrain map (1300 x 2100 int16, 10 x 20 sub-terrains), CUDA events, L2 flushed between calls, against the
HBM bound of the bytes written.  usage: python tools/terrain_bench.py [--cpu-reference]"""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

from pointfoot_b200 import capi, config
from pointfoot_b200.terrain import Terrain

dev = "cuda:0"
t0 = time.time()
t = Terrain(config.LeggedRobotCfg.terrain(), 4096, device=dev, seed=1)
torch.cuda.synchronize()
first = time.time() - t0
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = t._lib
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def timed(fn, n=20):
    tot = 0.0
    for i in range(n + 3):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        if i >= 3:
            tot += a.elapsed_time(b)
    return tot / n


gen = timed(lambda: capi.check(lib.pf_terrain_generate(C.byref(t._params), capi.ptr(t._subs), capi.ptr(t.height_field_raw),
                                                        capi.ptr(t.env_origins_device), st)))
R, Cn = t.height_field_raw.shape
tri = timed(lambda: capi.check(lib.pf_terrain_trimesh(capi.ptr(t.height_field_raw), C.c_int32(R), C.c_int32(Cn),
                                                       C.c_double(t.cfg.horizontal_scale), C.c_double(t.cfg.vertical_scale),
                                                       C.c_double(t.cfg.slope_treshold), capi.ptr(t.vertices), capi.ptr(t.triangles), st)))
peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))
except (OSError, ValueError):
    pass
hbm = float(peaks.get("hbm_gbs", peaks.get("hbm_gb_s", 6650.0))) if isinstance(peaks, dict) else 6650.0
gen_bytes = 2 * R * Cn
tri_bytes = 2 * R * Cn + 12 * R * Cn + 24 * (R - 1) * (Cn - 1)
out = {"map": [R, Cn], "first_construction_s_including_library_load": first,
       "pf_terrain_generate": {"us": gen * 1e3, "algorithmic_bytes": gen_bytes, "GB/s": gen_bytes / gen / 1e6, "frac_of_hbm": gen_bytes / gen / 1e6 / hbm},
       "pf_terrain_trimesh": {"us": tri * 1e3, "algorithmic_bytes": tri_bytes, "GB/s": tri_bytes / tri / 1e6, "frac_of_hbm": tri_bytes / tri / 1e6 / hbm},
       "hbm_peak_gbs": hbm}
print(json.dumps(out))
