"""CPU-side checks that need no GPU: the C-ABI library loads and exports every symbol that
include/ps_engine.h declares, refuses to run without a CUDA device (no CPU fallback), and the host
flattening accepts / rejects models as documented."""
import ctypes
import os
import re
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def englib():
    from mujoco_mpc_b200 import build
    build.build()
    from mujoco_mpc_b200 import engine
    return engine.lib()


def test_every_declared_symbol_is_exported(englib):
    hdr = open(os.path.join(ROOT, "include", "ps_engine.h")).read()
    names = set(re.findall(r"\b(ps_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    for n in sorted(names):
        assert hasattr(englib, n), f"{n} declared in include/ps_engine.h but not exported"


def test_no_cpu_fallback(englib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    h = englib.ps_create(0)
    assert not h
    assert b"no CUDA device" in englib.ps_last_error()
    from mujoco_mpc_b200 import registry
    from mujoco_mpc_b200.engine import Engine, EngineError
    m, t, s = registry.load("Cartpole")
    with pytest.raises(EngineError):
        Engine(m, t)


def test_model_header_in_sync_with_field_table():
    from mujoco_mpc_b200 import psmodel
    hdr = open(os.path.join(ROOT, "include", "ps_model.h")).read()
    for name, kind, shape in psmodel.FIELDS:
        assert re.search(rf"\b{name};", hdr), name


@pytest.mark.parametrize("name,nq,nv,nu,na", [("Cartpole", 2, 2, 1, 0), ("Acrobot", 2, 2, 1, 0), ("Particle", 2, 2, 2, 0),
                                              ("Swimmer", 8, 8, 5, 5), ("Walker", 9, 9, 6, 0)])
def test_task_models_compile(name, nq, nv, nu, na):
    from mujoco_mpc_b200 import registry
    m, t, s = registry.load(name)
    assert (m.nq, m.nv, m.nu, m.na) == (nq, nv, nu, na)
    assert t.num_residual == sum(t.dim_norm_residual[:t.num_term])
    assert s["steps"] == int(s["horizon"] / s["timestep"] + 1)
    M = np.asarray(m.dof_invweight0)
    assert np.all(M > 0)


def test_planner_library_exports_every_wrapper(englib):
    # the host planner classes (sampling / cross-entropy / robust) behind their extern "C" wrappers
    from mujoco_mpc_b200 import build, planner
    build.build_host()
    src = open(os.path.join(ROOT, "mujoco_mpc_b200", "host", "planner_capi.cc")).read()
    names = set(re.findall(r"^(?:void\*?|int|double)\s+(p(?:sp|ce|rb)_[a-z0-9_]+)\s*\(", src, flags=re.M))
    assert {"psp_create", "pce_create", "prb_create", "prb_optimize_policy", "prb_results", "psp_set_policy"} <= names
    L = planner.lib()
    for n in sorted(names):
        assert hasattr(L, n), n
