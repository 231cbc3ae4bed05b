"""Counter-based RNG contract, numpy statement.  ORACLE / TEST CODE.

torch's Philox offsets depend on how many envs reset in a step (point_foot.py:557, 577, 583, 632 draw
``(len(env_ids), k)`` tensors), so exact stream replay is impractical.  Kernels and oracle instead share
this contract: every random number is a pure function of

    (seed, step, purpose, global_env_id, lane)

``step`` is ``common_step_counter`` *after* its increment at point_foot.py:205.  The CUDA statement
of the same function lives in pointfoot_b200/csrc/pf_rng.cuh; tests compare the two bit for bit.
"""
import contextlib
import numpy as np
import torch

# purposes
OBS_NOISE = 0       # point_foot.py:332 (lanes 0..26) and :336 (lanes 27..PO-1)
RESET_DOF = 1       # point_foot.py:557, lanes 0..D-1
RESET_ROOT = 2      # point_foot.py:577 xy offset = lanes 0..1 (custom origins only); :583 velocities = lanes 2..7
CURRICULUM = 3      # point_foot.py:632 randint_like, lane 0
PUSH = 4            # point_foot.py:598, lanes 0..2
RESAMPLE_CMD = 5    # legged_robot.py:337-351 called from the callback (LR:326): lanes 0, 1, and 2 = the third draw (heading OR yaw)
RESAMPLE_RESET = 8  # the same method called from reset_idx (LR:166) on the same step: its own purpose, same lanes
TERRAIN = 9         # terrain generation (terrain.py + terrain_utils): 'env' = sub-terrain index, lanes = oracle/terrain_ref.py LANE_*; step 0
POLICY = 6          # rsl_rl ActorCritic.act: Normal.sample, lanes 0..A-1 (two u32 per lane)

_M = np.uint64(0xFFFFFFFF)


def _fmix32(h):
    h = h & _M
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x85EBCA6B)) & _M
    h ^= h >> np.uint64(13)
    h = (h * np.uint64(0xC2B2AE35)) & _M
    h ^= h >> np.uint64(16)
    return h


def rand_u32(seed, step, purpose, env, lane):
    """uint32 hash; env / lane broadcast as numpy arrays."""
    env = np.asarray(env, dtype=np.uint64)
    lane = np.asarray(lane, dtype=np.uint64)
    k = _fmix32(np.uint64((int(seed) * 0x9E3779B1 + int(step)) & 0xFFFFFFFF))
    k = _fmix32(k ^ np.uint64((int(purpose) * 0x85EBCA77) & 0xFFFFFFFF))
    k = _fmix32((k + env * np.uint64(0xC2B2AE3D)) & _M)
    k = _fmix32(k ^ ((lane * np.uint64(0x27D4EB2F)) & _M))
    return k.astype(np.uint32)


def uniform(seed, step, purpose, env, lane):
    """fp32 in [0, 1): top 24 bits * 2^-24 (same mapping torch.rand uses for float)."""
    u = rand_u32(seed, step, purpose, env, lane)
    return ((u >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)


def normal(seed, step, purpose, env, lane):
    """Box-Muller on lanes (2*lane, 2*lane+1): sqrt(-2 ln u1) * cos(2 pi u2), u1 in (0, 1]."""
    lane = np.asarray(lane, dtype=np.uint64)
    h1 = rand_u32(seed, step, purpose, env, 2 * lane)
    h2 = rand_u32(seed, step, purpose, env, 2 * lane + 1)
    u1 = ((h1 >> np.uint32(8)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)
    u2 = (h2 >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    r = np.sqrt(np.float32(-2.0) * np.log(u1, dtype=np.float32), dtype=np.float32)
    return (r * np.cos(np.float32(6.283185307179586) * u2, dtype=np.float32)).astype(np.float32)


# ---------------------------------------------------------------------------------------------
# Draw context: lets unmodified reference call sites (torch.rand_like, torch_rand_float,
# torch.randint_like) pull from the contract.  The harness wraps each method that draws.
class DrawContext:
    def __init__(self, seed, step, purpose, env_ids, env_offset=0, lane0=0):
        self.seed, self.step, self.purpose = seed, step, purpose
        self.env_ids = np.asarray(env_ids, dtype=np.int64) + int(env_offset)
        self.lane0 = lane0

    def uniform(self, shape):
        shape = tuple(int(s) for s in shape)
        n = shape[0]
        k = int(np.prod(shape[1:])) if len(shape) > 1 else 1
        assert n == len(self.env_ids), (shape, len(self.env_ids))
        lanes = np.arange(self.lane0, self.lane0 + k, dtype=np.int64)
        u = uniform(self.seed, self.step, self.purpose, self.env_ids[:, None], lanes[None, :])
        self.lane0 += k
        return u.reshape(shape)

    def randint(self, shape, high):
        shape = tuple(int(s) for s in shape)
        assert shape == (len(self.env_ids),)
        h = rand_u32(self.seed, self.step, self.purpose, self.env_ids, self.lane0)
        return (h.astype(np.int64) % int(high)).reshape(shape)


_active = []


@contextlib.contextmanager
def draw(ctx: DrawContext):
    _active.append(ctx)
    try:
        yield ctx
    finally:
        _active.pop()


def active_uniform(shape, device='cpu'):
    if not _active:
        return None
    return torch.from_numpy(_active[-1].uniform(shape)).to(device)


def active_randint(shape, high, device='cpu'):
    if not _active:
        return None
    return torch.from_numpy(_active[-1].randint(shape, high)).to(device)
