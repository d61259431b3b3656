"""The reference-side binding (mujoco_mpc_b200/host/mjpc_adapter: mjpc::CudaSamplingAdapter : public mjpc::RankedPlanner,
compiled against the reference's own planner.h / task.h / state.h) driven ONLY through mjpc::Planner* virtuals with the
reference's Task, State and ThreadPool objects -- next to the reference's own mjpc::SamplingPlanner driven by the very same
call sequence.  CPU: the adapter runs on the host stand-in of the engine ABI; GPU: on libps_engine.so."""
import ctypes
import numpy as np
import pytest
from mujoco_mpc_b200 import registry
from tests.adapter import build as adapter_build
from tests.common import model_mocap

_dp = ctypes.POINTER(ctypes.c_double)


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


class Harness:
    def __init__(self, lib, model, name, kind):
        self.L, self.model = lib, model
        lib.adp_create.restype = ctypes.c_void_p
        self._keep = model.c_struct()
        self.h = ctypes.c_void_p(lib.adp_create(ctypes.byref(self._keep), name.encode(), kind, 0))
        assert self.h, name
        self.ns = model.nq + model.nv + model.na

    def close(self):
        if self.h:
            self.L.adp_destroy(self.h); self.h = None

    def configure(self, P, interp, s0, s1, N):
        self.L.adp_configure(self.h, P, interp, ctypes.c_double(s0), ctypes.c_double(s1), N)

    def plan(self, state, mocap, time, z, N, P, H):
        s = np.ascontiguousarray(state, dtype=np.float64); zz = np.ascontiguousarray(z, dtype=np.float64)
        mo = np.ascontiguousarray(mocap, dtype=np.float64) if mocap is not None else None
        self.L.adp_plan(self.h, _p(s), _p(mo), ctypes.c_double(time), _p(zz), None, N, P, H)
        self.H = H

    def action(self, t, prev=False):
        out = np.zeros(max(self.model.nu, 1)); self.L.adp_action(self.h, ctypes.c_double(t), int(prev), _p(out)); return out[:self.model.nu]

    def best(self, nr):
        H = self.H
        st = np.zeros((H, self.ns)); ac = np.zeros((H, self.model.nu)); tm = np.zeros(H); rs = np.zeros((H, nr)); cs = np.zeros(H)
        ret = ctypes.c_double()
        h = self.L.adp_best(self.h, _p(st), _p(ac), _p(tm), _p(rs), _p(cs), ctypes.byref(ret))
        assert h == H
        return dict(states=st, actions=ac, times=tm, residual=rs, costs=cs, total_return=ret.value)


def run_pair(lib, name, interp, iters=3):
    model, task, s = registry.load(name)
    N, H, P = 24, min(s["steps"], 51), min(s["num_spline_points"], 36)
    ours = Harness(lib, model, name, 0); ref = Harness(lib, model, name, 1)
    for h in (ours, ref):
        h.configure(P, interp, s["exploration"], 0.0, N)
    assert ours.L.adp_num_parameters(ours.h) == ref.L.adp_num_parameters(ref.h) == P * model.nu
    rng = np.random.default_rng(5)
    state = registry.synthetic_state(model); t = 0.05; mocap = model_mocap(model)
    for it in range(iters):
        z = rng.standard_normal((N, P, model.nu))
        ours.plan(state, mocap, t, z, N, P, H); ref.plan(state, mocap, t, z, N, P, H)
        for tq in (t, t + 0.017, t + 0.3, t + 9.0):
            assert np.allclose(ours.action(tq), ref.action(tq), rtol=0, atol=1e-13), (name, it, tq)
            assert np.allclose(ours.action(tq, True), ref.action(tq, True), rtol=0, atol=1e-13), (name, it, tq)
        a, b = ours.best(task.num_residual), ref.best(task.num_residual)
        assert np.abs(a["states"] - b["states"]).max() <= 1e-9 * max(1.0, np.abs(b["states"]).max()), (name, it)
        assert np.allclose(a["actions"], b["actions"], rtol=0, atol=1e-12)
        assert np.abs(a["residual"] - b["residual"]).max() <= 1e-8 * max(1.0, np.abs(b["residual"]).max())
        assert abs(a["total_return"] - b["total_return"]) <= 1e-8 * max(1.0, abs(b["total_return"]))
        state = b["states"][2].copy(); t = float(b["times"][2])
    ours.close(); ref.close()


@pytest.mark.parametrize("name,interp", [("Cartpole", 2), ("Particle", 1), ("Acrobot", 0), ("Walker", 2)])
def test_adapter_behind_the_planner_interface_matches_the_reference_planner_cpu(oracle_lib, name, interp):
    path = adapter_build.build("host")
    if path is None:
        pytest.skip("tests/adapter/libadapter_host.so can only be built where /root/reference exists")
    run_pair(ctypes.CDLL(path), name, interp)


@pytest.mark.gpu
@pytest.mark.parametrize("name,interp", [("Cartpole", 2), ("Particle", 1), ("Swimmer", 2), ("Walker", 2)])
def test_adapter_behind_the_planner_interface_matches_the_reference_planner_gpu(name, interp):
    path = adapter_build.build("cuda")
    if path is None:
        pytest.skip("tests/adapter/libadapter_cuda.so was not built (needs /root/reference at build time)")
    run_pair(ctypes.CDLL(path), name, interp)
