"""The C-ABI library loads without a GPU and exports every symbol include/pointfoot_b200.h declares; the ctypes
struct layouts match the header; compute entry points fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import os
import re

import pytest
import torch

import common
from pointfoot_b200 import capi, synth

HEADER = os.path.join(common.ROOT, "include", "pointfoot_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pf_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.isfile(capi.lib_path()):
        import __graft_entry__
        __graft_entry__.build()
    return capi.load()


def test_exports_every_declared_symbol(lib):
    names = declared_functions()
    assert "pf_post_physics_step" in names and "pf_pd_torques" in names
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"


def test_struct_layouts_match(lib):
    assert lib.pf_abi_version() == capi.PF_ABI_VERSION
    assert lib.pf_sizeof_params() == C.sizeof(capi.PfParams)
    assert lib.pf_sizeof_env_buffers() == C.sizeof(capi.PfEnvBuffers)


def test_no_cpu_fallback(lib):
    if torch.cuda.is_available():
        pytest.skip("has a device")
    st = synth.make_state(8, "pointfoot_flat")
    cfg, P, noise, pts, names = common.make_params("pointfoot_flat", st)
    ctx = C.c_void_p()
    rc = lib.pf_ctx_create(C.byref(P), C.c_void_p(noise.data_ptr()), None, C.byref(ctx))
    assert rc != 0 and b"no CUDA device" in lib.pf_last_error()


def test_bad_params_are_rejected(lib):
    st = synth.make_state(8, "pointfoot_flat")
    cfg, P, noise, pts, names = common.make_params("pointfoot_flat", st)
    P.abi_version = 99
    ctx = C.c_void_p()
    assert lib.pf_ctx_create(C.byref(P), C.c_void_p(noise.data_ptr()), None, C.byref(ctx)) != 0
    cfg, P, noise, pts, names = common.make_params("pointfoot_flat", st)
    P.num_obs = 30            # reference: RuntimeError at point_foot.py:363-365
    rc = lib.pf_ctx_create(C.byref(P), C.c_void_p(noise.data_ptr()), None, C.byref(ctx))
    assert rc == 3 and b"observation width" in lib.pf_last_error()


def test_redzone_helper_detects_overruns_on_either_side():
    """tests/common.py::RedZone is the memcheck stand-in of tests/test_redzone_gpu.py: it must flag a byte written before or
    after a wrapped buffer and accept writes inside it"""
    import pytest
    import torch
    import common
    rz = common.RedZone()
    t = rz.wrap(torch.arange(10, dtype=torch.float32), "t")
    t += 1
    assert rz.check() == 1 and float(t[9]) == 10.0
    name, raw, n = rz.guards[0]
    raw[rz.PAD + n] = 0
    with pytest.raises(AssertionError, match="ABOVE buffer t"):
        rz.check()
    raw[rz.PAD + n] = 0xA5
    raw[rz.PAD - 1] = 1
    with pytest.raises(AssertionError, match="BELOW buffer t"):
        rz.check()


def test_bench_sweep_cases_cover_c5_once_per_case():
    """bench.py's size sweep: nine single-GPU cases; under torchrun the 262,144-env job sharded over the ranks plus 32,768
    envs per GPU -- which coincide at 8 ranks and must then run once (host logic only: nothing is launched)"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert len(bench.sweep_cases(1)) == 9 and ("pointfoot_rough", 262144, True) in bench.sweep_cases(1)
    for world in (2, 4, 8):
        cases = bench.sweep_cases(world)
        assert len(cases) == len(set(cases))
        assert {n for _, n, _ in cases} == {262144 // world, 32768}
        assert {t for t, _, _ in cases} == {"pointfoot_flat", "pointfoot_rough"}
