"""Terrain generation on the GPU through the C-ABI (pf_terrain_generate / pf_terrain_trimesh):
against golden maps of the UNMODIFIED reference Terrain class, against the host emulation at another seed, and as the
height field of the step path."""
import numpy as np
import pytest
import torch

import parity_terrain

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fname", list(parity_terrain.TERRAIN_GOLDEN))
def test_cuda_raster_matches_reference_golden(fname):
    parity_terrain.golden_impl(fname, "cuda:0")


@pytest.mark.parametrize("mutate", [parity_terrain.t_all_kinds_random, parity_terrain.t_all_kinds_curriculum])
def test_cuda_raster_matches_host_emulation_at_another_seed(mutate):
    from pointfoot_b200.terrain import Terrain
    a = Terrain(parity_terrain.make_cfg(mutate), 64, device="cuda:0", seed=77)
    b = parity_terrain.emu_class()(parity_terrain.make_cfg(mutate), 64, device="cpu", seed=77)
    assert torch.equal(a.height_field_raw.cpu(), b.height_field_raw)
    assert np.array_equal(a.env_origins, b.env_origins)
    assert torch.equal(a.vertices.cpu(), b.vertices) and torch.equal(a.triangles.cpu(), b.triangles)
    c = Terrain(parity_terrain.make_cfg(mutate), 64, device="cuda:0", seed=78)
    assert not torch.equal(a.height_field_raw, c.height_field_raw)


def test_generated_field_feeds_the_step_path():
    """the map the raster kernel leaves in HBM is the int16 field the height scan gathers from (no copy, no host trip)"""
    from pointfoot_b200 import config, synth
    from pointfoot_b200.legged_robot import LeggedRobot
    from pointfoot_b200.terrain import Terrain
    t = Terrain(config.LeggedRobotCfg.terrain(), 256, device="cuda:0", seed=1)
    state = synth.make_lr_state(256, seed=2)
    state["height_samples"] = t.heightsamples
    state["terrain_origins"] = torch.from_numpy(t.env_origins).to(torch.float)
    cfg = config.A1RoughCfg()
    cfg.env.num_envs = 256
    env = LeggedRobot(cfg, None, None, "cuda:0", True, state=state)
    assert env.height_samples.data_ptr() == t.heightsamples.data_ptr()
    obs, _, rew, reset, _ = env.step(torch.zeros(256, 12, device="cuda"))
    assert torch.isfinite(obs).all() and float(env.measured_heights.abs().max()) > 0.05      # real relief under the robots
    # the scan's samples are values of the generated map times the vertical scale
    vals = torch.unique(t.heightsamples).float() * cfg.terrain.vertical_scale
    assert torch.isin(env.measured_heights.flatten(), vals).all()
