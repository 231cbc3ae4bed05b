"""Terrain generation, CPU suite: the per-pixel code of the raster kernels (pf_terrain_core.cuh,
driven by tests/host_emu) against golden maps of the UNMODIFIED reference Terrain class; the reference re-run live where it
exists; the one piece of the restated third-party module that CAN be cross-checked here (the interp2d replacement)."""
import numpy as np
import pytest

import parity_terrain
from oracle import ref_harness, terrain_ref

needs_reference = pytest.mark.skipif(not ref_harness.reference_available(), reason="/root/reference not present")


@pytest.mark.parametrize("fname", list(parity_terrain.TERRAIN_GOLDEN))
def test_raster_core_host_emulation_matches_reference_golden(fname):
    parity_terrain.golden_impl(fname, "cpu")


@needs_reference
@pytest.mark.parametrize("fname", ["terrain_all_random.npz", "terrain_all_curriculum.npz"])
def test_live_reference_reproduces_golden(fname):
    z = np.load(parity_terrain.os.path.join(parity_terrain.GOLDEN_DIR, fname))
    t = ref_harness.build_reference_terrain(parity_terrain.TERRAIN_GOLDEN[fname][0], seed=1)
    assert np.array_equal(t.height_field_raw, z["height_field"]) and np.array_equal(t.env_origins, z["env_origins"])


def test_every_make_terrain_branch_is_in_the_golden_maps():
    t = parity_terrain.make_impl("terrain_all_random.npz", "cpu")
    kinds = set(int(k) for k in t._subs[:, 0])
    t2 = parity_terrain.make_impl("terrain_all_curriculum.npz", "cpu")
    assert kinds | set(int(k) for k in t2._subs[:, 0]) == set(range(7))


def test_bilinear_restatement_equals_scipys_interp2d_replacement():
    """random_uniform_terrain calls scipy.interpolate.interp2d(kind='linear') (removed in SciPy 1.14); SciPy names
    RectBivariateSpline as its replacement on regular grids.  The two agree to 1e-12 before np.rint and on every pixel after
    it wherever an exact .5 cannot occur -- which is every shape the reference produces: W px upsampled from W / 2 nodes gives
    fractions with the odd denominator W - 1 (79 for the 8 m cell, 59 for 6 m).  (7 -> 33 has ties: FITPACK's rounding noise
    decides those, so only the 1e-12 bar is asserted there.)"""
    from scipy.interpolate import RectBivariateSpline
    g = np.random.default_rng(0)
    for n_in, n_out in ((40, 80), (30, 60), (7, 33)):
        grid = g.integers(-10, 11, (n_in, n_in)).astype(np.float64)
        x = np.linspace(0, 8.0, n_in)
        xu = np.linspace(0, 8.0, n_out)
        want = RectBivariateSpline(x, x, grid, kx=1, ky=1)(xu, xu)
        got = terrain_ref.bilinear_upsample(grid, n_out, n_out)
        assert np.abs(got - want).max() < 1e-12
        if (n_out - 1) % 2 == 1:
            assert np.array_equal(np.rint(got), np.rint(want))


def test_terrain_has_no_cpu_fallback_and_mirrors_the_reference_errors():
    from pointfoot_b200 import config
    from pointfoot_b200.terrain import Terrain
    cfg = config.LeggedRobotCfg.terrain()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Terrain(cfg, 16, device="cpu")
    cfg = config.LeggedRobotCfg.terrain()
    cfg.mesh_type = "plane"
    assert not hasattr(Terrain(cfg, 16, device="cpu"), "height_field_raw")       # terrain.py:44-45 returns early
    emu = parity_terrain.emu_class()
    cfg = config.LeggedRobotCfg.terrain()
    cfg.curriculum, cfg.selected, cfg.terrain_kwargs = False, True, {"type": "terrain_utils.wave_terrain"}
    with pytest.raises(AttributeError, match="vertical_scale"):                  # terrain.py:100-103
        emu(cfg, 16, device="cpu")
    cfg = config.LeggedRobotCfg.terrain()
    cfg.terrain_proportions = [0.1, 0.1, 0.3, 0.2, 0.1]                           # sums to 0.8: the last columns fall off the list
    with pytest.raises(IndexError):                                              # terrain.py:139 proportions[5]
        emu(cfg, 16, device="cpu")
    cfg = config.LeggedRobotCfg.terrain()
    cfg.terrain_length = 6.
    with pytest.raises(ValueError, match="could not broadcast"):                 # terrain.py:155
        emu(cfg, 16, device="cpu")
