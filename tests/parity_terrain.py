"""Parity driver of terrain generation, SURVEY.md section 8f row N1.  TEST INFRASTRUCTURE.

The checker is the UNMODIFIED legged_gym/utils/terrain.py::Terrain run over the restated isaacgym.terrain_utils
(oracle/terrain_ref.py: third-party module absent from the reference tree -> PARITY UNPINNED for the five primitives it
restates; the reference's own layout, paste, env-origin, gap and pit code is pinned).  Outputs are committed as
tests/golden/terrain_*.npz (tests/golden/make_golden.py::generate_terrain).  Bars: height field int16 BIT-EXACT, env_origins
float64 bit-exact, trimesh vertices fp32 bit-exact, triangles exact."""
import ctypes as C
import hashlib
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def t_default(cfg):
    """LeggedRobotCfg.terrain as it stands: curriculum, 10 x 20 cells of 8 m, 25 m border, 5 proportions (types 0-3 reachable)"""


def t_all_kinds_random(cfg):
    """every branch of make_terrain (8 cumulative proportions), randomized layout (choice / difficulty drawn), small border"""
    cfg.curriculum = False
    cfg.terrain_proportions = [0.1, 0.1, 0.2, 0.15, 0.15, 0.1, 0.1, 0.1]
    cfg.num_rows, cfg.num_cols, cfg.border_size = 4, 7, 2


def t_all_kinds_curriculum(cfg):
    """every branch, curriculum layout (difficulty 0 included: stone distance 0.05 -> 0 px), other scales and cell size"""
    cfg.terrain_proportions = [0.15, 0.1, 0.15, 0.15, 0.15, 0.1, 0.1, 0.1]
    cfg.num_rows, cfg.num_cols, cfg.border_size = 5, 10, 1.5
    cfg.terrain_length = cfg.terrain_width = 6.
    cfg.vertical_scale = 0.004
    cfg.slope_treshold = 0.5


# file -> (mutation, keep the whole field in the file?)  (the default map is 5.5 MB: its sha256 and a 160-row slab are kept)
TERRAIN_GOLDEN = {
    "terrain_default.npz": (t_default, False),
    "terrain_all_random.npz": (t_all_kinds_random, True),
    "terrain_all_curriculum.npz": (t_all_kinds_curriculum, True),
}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def make_cfg(mutate):
    from pointfoot_b200 import config
    cfg = config.LeggedRobotCfg.terrain()
    mutate(cfg)
    return cfg


def emu_class():
    """Terrain with its two C-ABI calls redirected to the host emulation of the same per-pixel code (tests/host_emu)"""
    import common
    from pointfoot_b200 import capi
    from pointfoot_b200.terrain import Terrain

    class EmuTerrain(Terrain):
        def _open_backend(self):
            self._lib = common.host_emu()

        def _call_generate(self, P, subs, hf, origins):
            assert self._lib.pf_emu_terrain_generate(C.byref(P), capi.ptr(subs), capi.ptr(hf), capi.ptr(origins)) == 0

        def _call_trimesh(self, hf, R, Cn, hs, vs, thr, vertices, triangles):
            assert self._lib.pf_emu_terrain_trimesh(capi.ptr(hf), C.c_int(R), C.c_int(Cn), C.c_double(hs), C.c_double(vs),
                                                    C.c_double(thr), capi.ptr(vertices), capi.ptr(triangles)) == 0
    return EmuTerrain


def make_impl(fname, device="cuda:0"):
    from pointfoot_b200.terrain import Terrain
    cls = emu_class() if device == "cpu" else Terrain
    return cls(make_cfg(TERRAIN_GOLDEN[fname][0]), 64, device=device, seed=1)


def check_against_golden(fname, t):
    z = np.load(os.path.join(GOLDEN_DIR, fname))
    hf = t.height_field_raw.cpu().numpy()
    assert hf.dtype == np.int16 and hf.shape == tuple(z["shape"])
    assert (t.tot_rows, t.tot_cols, t.border, t.width_per_env_pixels) == tuple(int(v) for v in z["geometry"])
    if "height_field" in z.files:
        bad = hf != z["height_field"]
        assert not bad.any(), f"[{fname}] {int(bad.sum())} of {bad.size} pixels differ from the reference (first at {np.argwhere(bad)[0]})"
    else:
        r0, r1 = (int(v) for v in z["slab_rows"])
        assert np.array_equal(hf[r0:r1], z["slab"]), f"[{fname}] slab rows {r0}:{r1} differ from the reference"
    assert sha(hf) == bytes(z["hf_sha256"]).decode(), f"[{fname}] height field checksum"
    assert np.array_equal(t.env_origins, z["env_origins"]), f"[{fname}] env_origins"
    assert t.env_origins.dtype == np.float64
    if t.type == "trimesh":
        v, tri = t.vertices.cpu().numpy(), t.triangles.cpu().numpy().view(np.uint32)
        assert sha(v) == bytes(z["vertices_sha256"]).decode(), f"[{fname}] trimesh vertices"
        assert sha(tri) == bytes(z["triangles_sha256"]).decode(), f"[{fname}] trimesh triangles"
        assert int(z["moved_vertices"]) > 0


def golden_impl(fname, device="cuda:0"):
    check_against_golden(fname, make_impl(fname, device))
