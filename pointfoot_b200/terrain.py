"""``Terrain`` -- drop-in for legged_gym/utils/terrain.py::Terrain (terrain.py:38-187), SURVEY.md section 8f row N1.

Same constructor ``Terrain(cfg, num_robots)`` and attributes (``env_length``, ``env_width``, ``proportions``, ``env_origins``,
``width_per_env_pixels``, ``length_per_env_pixels``, ``border``, ``tot_rows``, ``tot_cols``, ``height_field_raw``,
``heightsamples``, ``vertices``, ``triangles``), but the map is rasterised on the GPU by ``pf_terrain_generate`` (one thread
per pixel of the whole field, include/pointfoot_b200.h) instead of 200 numpy sub-terrains pasted into it: ``height_field_raw``
/ ``heightsamples`` are CUDA int16 tensors the step path's height scan reads in place; ``env_origins`` is copied back as the
float64 numpy array the reference keeps (600 numbers, init time) and kept on the device as ``env_origins_device``.

The host part below is the reference's own parameter arithmetic (``make_terrain``, terrain.py:109-145) followed by the
integer reductions the isaacgym.terrain_utils primitives apply to their arguments (``int(step_height / vertical_scale)``
...), in Python doubles like the reference; random structure (noise grid, rectangles, stone offsets, and the per-cell
choice / difficulty of ``randomized_terrain``) follows the counter-based contract of csrc/pf_rng.cuh, purpose TERRAIN.
No CPU fallback: without the CUDA library or a device the constructor raises.
"""
import ctypes as C

import numpy as np
import torch

from . import capi

KIND = {"slope": 0, "noisy_slope": 1, "stairs": 2, "obstacles": 3, "stones": 4, "gap": 5, "pit": 6}
RNG_TERRAIN, LANE_LAYOUT = 9, 0x10000


def _fmix32(h):
    h &= 0xFFFFFFFF
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & 0xFFFFFFFF
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & 0xFFFFFFFF
    h ^= h >> 16
    return h


def contract_u32(seed, sub, lane, step=0, purpose=RNG_TERRAIN):
    """csrc/pf_rng.cuh: pf_rng_u32(pf_rng_env_key(pf_rng_key(seed, step, purpose), sub), lane)"""
    k = _fmix32((seed * 0x9E3779B1 + step) & 0xFFFFFFFF)
    k = _fmix32(k ^ ((purpose * 0x85EBCA77) & 0xFFFFFFFF))
    k = _fmix32((k + sub * 0xC2B2AE3D) & 0xFFFFFFFF)
    return _fmix32(k ^ ((lane * 0x27D4EB2F) & 0xFFFFFFFF))


def _rect(a, b, c, d, n):
    """numpy's reading of field[a:b, c:d] on an n x n array (negative indices wrap, ends clamp) -> x1 x2 y1 y2"""
    x1, x2, _ = slice(a, b).indices(n)
    y1, y2, _ = slice(c, d).indices(n)
    return [x1, max(x1, x2), y1, max(y1, y2)]


class Terrain:
    def __init__(self, cfg, num_robots, device="cuda:0", seed=1):
        self.cfg = cfg
        self.num_robots = num_robots
        self.type = cfg.mesh_type
        if self.type in ["none", "plane"]:
            return
        self.device = torch.device(device)
        self._open_backend()
        self.seed = int(seed)
        self.env_length = cfg.terrain_length
        self.env_width = cfg.terrain_width
        self.proportions = [np.sum(cfg.terrain_proportions[:i + 1]) for i in range(len(cfg.terrain_proportions))]
        self.cfg.num_sub_terrains = cfg.num_rows * cfg.num_cols
        self.width_per_env_pixels = int(self.env_width / cfg.horizontal_scale)
        self.length_per_env_pixels = int(self.env_length / cfg.horizontal_scale)
        self.border = int(cfg.border_size / self.cfg.horizontal_scale)
        self.tot_cols = int(cfg.num_cols * self.width_per_env_pixels) + 2 * self.border
        self.tot_rows = int(cfg.num_rows * self.length_per_env_pixels) + 2 * self.border
        if self.length_per_env_pixels != self.width_per_env_pixels:
            # terrain.py:110-114 builds every sub-terrain width x width and pastes it into a length x width cell (:155)
            raise ValueError(f"could not broadcast input array from shape ({self.width_per_env_pixels},{self.width_per_env_pixels}) "
                             f"into shape ({self.length_per_env_pixels},{self.width_per_env_pixels})")
        if cfg.curriculum:
            subs = self.curiculum()
        elif cfg.selected:
            subs = self.selected_terrain()
        else:
            subs = self.randomized_terrain()
        self._generate(subs)
        self.heightsamples = self.height_field_raw
        if self.type == "trimesh":
            self.vertices, self.triangles = self.convert_heightfield_to_trimesh(
                self.height_field_raw, self.cfg.horizontal_scale, self.cfg.vertical_scale, self.cfg.slope_treshold)

    # ------------------------------------------------------------------ layout (terrain.py:73-107): one record per map cell
    def randomized_terrain(self):
        subs = [None] * self.cfg.num_sub_terrains
        for k in range(self.cfg.num_sub_terrains):
            # np.random.uniform(0, 1) / np.random.choice([0.5, 0.75, 0.9]) of terrain.py:78-79, from the contract
            u = np.float32(contract_u32(self.seed, k, LANE_LAYOUT) >> 8) * np.float32(1.0 / 16777216.0)
            choice = 0.0 + (1.0 - 0.0) * float(u)
            difficulty = [0.5, 0.75, 0.9][contract_u32(self.seed, k, LANE_LAYOUT + 1) % 3]
            subs[k] = self.make_terrain(choice, difficulty)
        return subs

    def curiculum(self):
        subs = [None] * self.cfg.num_sub_terrains
        for j in range(self.cfg.num_cols):
            for i in range(self.cfg.num_rows):
                difficulty = i / self.cfg.num_rows
                choice = j / self.cfg.num_cols + 0.001
                subs[i * self.cfg.num_cols + j] = self.make_terrain(choice, difficulty)
        return subs

    def selected_terrain(self):
        self.cfg.terrain_kwargs.pop('type')
        # terrain.py:100-103 reads self.vertical_scale / self.horizontal_scale, which Terrain never sets
        raise AttributeError("'Terrain' object has no attribute 'vertical_scale'")

    # ------------------------------------------------------------------ terrain.py:109-145 + the primitives' integer reductions
    def make_terrain(self, choice, difficulty):
        W, hs, vs = self.width_per_env_pixels, self.cfg.horizontal_scale, self.cfg.vertical_scale
        slope = difficulty * 0.4
        step_height = 0.05 + 0.18 * difficulty
        discrete_obstacles_height = 0.05 + difficulty * 0.2
        stepping_stones_size = 1.5 * (1.05 - difficulty)
        stone_distance = 0.05 if difficulty == 0 else 0.1
        gap_size = 1. * difficulty
        pit_depth = 1. * difficulty
        pr = self.proportions
        p = [0] * 15

        def platform_rect(platform_size):               # x1 = (width - platform) // 2, x2 = (width + platform) // 2 (both axes)
            ps = int(platform_size / hs)
            return _rect((W - ps) // 2, (W + ps) // 2, (W - ps) // 2, (W + ps) // 2, W)

        if choice < pr[1]:
            noisy = not choice < pr[0]
            if not noisy and choice < pr[0] / 2:
                slope *= -1
            p[0] = int(slope * (hs / vs) * (W / 2))                         # pyramid_sloped_terrain(slope, platform_size=3.)
            p[1] = int(3. / hs / 2)
            if not 0 <= W // 2 - p[1] < W:
                raise IndexError("pyramid_sloped_terrain: platform larger than the sub-terrain")
            if noisy:                                                       # random_uniform_terrain(-0.05, 0.05, 0.005, 0.2)
                lo, hi, st = int(-0.05 / vs), int(0.05 / vs), int(0.005 / vs)
                if st <= 0:
                    raise ZeroDivisionError("random_uniform_terrain: step below the vertical scale")
                p[2], p[3], p[4] = lo, len(range(lo, hi + st, st)), st
                p[5] = p[6] = int(W * hs / 0.2)
                if p[5] < 2 or p[3] < 1:
                    raise ValueError("random_uniform_terrain: degenerate coarse grid")
            return (KIND["noisy_slope" if noisy else "slope"], p)
        if choice < pr[3]:
            if choice < pr[2]:
                step_height *= -1
            sw, sh, ps = int(0.31 / hs), int(step_height / vs), int(3. / hs)  # pyramid_stairs_terrain(0.31, step_height, 3.)
            if sw <= 0:
                raise ValueError("pyramid_stairs_terrain: step width below the horizontal scale (the reference never terminates)")
            steps, start, stop = 0, 0, W
            while (stop - start) > ps:
                start, stop, steps = start + sw, stop - sw, steps + 1
            p[0], p[1], p[2] = sw, sh, steps
            return (KIND["stairs"], p)
        if choice < pr[4]:                                                   # discrete_obstacles_terrain(h, 1., 2., 20, 3.)
            p[0], p[1], p[2], p[3] = int(discrete_obstacles_height / vs), int(1. / hs), int(2. / hs), 20
            if len(range(p[1], p[2], 4)) == 0 or W - (p[2] - 1) <= 0:
                raise ValueError("discrete_obstacles_terrain: empty size / position range")
            p[4:8] = platform_rect(3.)
            return (KIND["obstacles"], p)
        if choice < pr[5]:                                                   # stepping_stones_terrain(size, distance, 0., 4.)
            p[0], p[1], p[2], p[3] = int(stepping_stones_size / hs), int(stone_distance / hs), int(0. / vs), int(-10 / vs)
            if p[0] <= 0:
                raise ValueError("stepping_stones_terrain: stone size below the horizontal scale")
            if (W + p[0] + p[1] - 1) // (p[0] + p[1]) > 126:
                raise ValueError("stepping_stones_terrain: more stones per band than the RNG lane layout holds")
            p[4:8] = platform_rect(4.)
            return (KIND["stones"], p)
        if choice < pr[6]:                                                   # gap_terrain(gap_size, 3.), terrain.py:169-181
            g, ps = int(gap_size / hs), int(3. / hs)
            c = W // 2
            x1 = (W - ps) // 2
            x2 = x1 + g
            p[0:4] = _rect(c - x2, c + x2, c - x2, c + x2, W)
            p[4:8] = _rect(c - x1, c + x1, c - x1, c + x1, W)
            return (KIND["gap"], p)
        depth, ps = int(pit_depth / vs), int(4. / hs / 2)                     # pit_terrain(depth, 4.), terrain.py:183-190
        p[0] = -depth
        p[4:8] = _rect(W // 2 - ps, W // 2 + ps, W // 2 - ps, W // 2 + ps, W)
        return (KIND["pit"], p)

    # ------------------------------------------------------------------ the launches
    def _open_backend(self):
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("pointfoot_b200.Terrain needs a CUDA device (no CPU fallback); got " + str(self.device))
        self._lib = capi.load_terrain()

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _call_generate(self, P, subs, hf, origins):
        capi.check(self._lib.pf_terrain_generate(C.byref(P), capi.ptr(subs), capi.ptr(hf), capi.ptr(origins), self._stream()),
                   "pf_terrain_generate")

    def _call_trimesh(self, hf, R, Cn, hs, vs, thr, vertices, triangles):
        capi.check(self._lib.pf_terrain_trimesh(capi.ptr(hf), C.c_int32(R), C.c_int32(Cn), C.c_double(hs), C.c_double(vs), C.c_double(thr),
                                                capi.ptr(vertices), capi.ptr(triangles), self._stream()), "pf_terrain_trimesh")

    def _generate(self, subs):
        cfg, W = self.cfg, self.width_per_env_pixels
        P = capi.PfTerrainParams()
        P.abi_version, P.num_rows, P.num_cols, P.cell_px = capi.PF_ABI_VERSION, cfg.num_rows, cfg.num_cols, W
        P.border_px, P.tot_rows, P.tot_cols, P.seed = self.border, self.tot_rows, self.tot_cols, self.seed & 0xFFFFFFFF
        hs = cfg.horizontal_scale                                           # terrain.py:162-165
        patch = _rect(int((self.env_length / 2. - 1) / hs), int((self.env_length / 2. + 1) / hs),
                      int((self.env_width / 2. - 1) / hs), int((self.env_width / 2. + 1) / hs), W)
        for k in range(4):
            P.origin_patch[k] = patch[k]
        P.env_length, P.env_width, P.vertical_scale = self.env_length, self.env_width, cfg.vertical_scale
        table = np.zeros((len(subs), 16), np.int32)
        for k, (kind, p) in enumerate(subs):
            table[k, 0], table[k, 1:] = kind, p
        self._subs = torch.from_numpy(table).to(self.device)
        self.height_field_raw = torch.empty(self.tot_rows, self.tot_cols, dtype=torch.int16, device=self.device)
        self.env_origins_device = torch.empty(cfg.num_rows, cfg.num_cols, 3, dtype=torch.float64, device=self.device)
        self._params = P
        self._call_generate(P, self._subs, self.height_field_raw, self.env_origins_device)
        self.env_origins = self.env_origins_device.cpu().numpy()

    def convert_heightfield_to_trimesh(self, height_field_raw, horizontal_scale, vertical_scale, slope_threshold=None):
        """isaacgym.terrain_utils.convert_heightfield_to_trimesh on the device: (vertices fp32[R*C, 3], triangles uint32[2(R-1)(C-1), 3])"""
        hf = height_field_raw.contiguous()
        R, Cn = hf.shape
        vertices = torch.empty(R * Cn, 3, dtype=torch.float32, device=hf.device)
        triangles = torch.empty(2 * (R - 1) * (Cn - 1), 3, dtype=torch.int32, device=hf.device)     # bit pattern of uint32
        self._call_trimesh(hf, R, Cn, horizontal_scale, vertical_scale, -1.0 if slope_threshold is None else slope_threshold,
                           vertices, triangles)
        return vertices, triangles
