"""Pins the planner / rollout / cost / residual side to the REFERENCE'S OWN CODE.

tests/golden/planner_reference.npz was written by tools/gen_golden_planner.py from mjpc::SamplingPlanner, SamplingPolicy,
Trajectory::NoisyRollout, Task / BaseResidualFn and the six tasks/*/*.cc compiled unmodified from the reference tree
(oracle/_ref/libmjpc_ref.so; physics below mj_step = the oracle's restated step, so the PHYSICS stays unpinned).  Checked
here, on the same inputs and the same injected knot noise:
  * CPU: the oracle's restatement (mm_planner.cc / mm_cost.cc / mm_quadruped.cc) reproduces the fixtures -- returns, winner,
    resampled policies, trajectories -- over two consecutive planning iterations, for all six tasks and the five quadruped modes;
  * GPU: the CUDA engine behind the drop-in host class does the same (north_star tolerances);
  * live (only where oracle/_ref is built, i.e. where /root/reference exists): fresh seeds beyond the fixtures, and the
    measured impact of the one semantic deviation of the engine (zero warm start per rollout).
"""
import os
import numpy as np
import pytest
from mujoco_mpc_b200 import registry
from oracle import pyoracle

GOLD = os.path.join(os.path.dirname(__file__), "golden", "planner_reference.npz")
STATE_RTOL, COST_RTOL = 1e-9, 1e-8


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD, allow_pickle=False)


def case_ids():
    return [str(c) for c in np.load(GOLD, allow_pickle=False)["cases"]]


def load_case(g, cid):
    """(model, task with the fixture's parameters / weights / residual state, settings, inputs)"""
    inp = {k.split("/in/")[1]: g[k] for k in g.files if k.startswith(cid + "/in/")}
    out = {k.split("/out/")[1]: g[k] for k in g.files if k.startswith(cid + "/out/")}
    model, task, s = registry.load(str(inp["task"]))
    for k, v in enumerate(inp["parameters"]):
        task.parameters[k] = float(v)
    for k, v in enumerate(inp["weight"]):
        task.weight[k] = float(v)
    for k, v in enumerate(inp["aux"]):
        task.aux[k] = float(v)
    return model, task, s, inp, out


def rel(a, b):
    return np.abs(a - b).max() / max(1.0, np.abs(b).max()) if a.size else 0.0


def check_iteration(cid, it, out, winner, returns, failure, pol, prev, traj_of):
    assert np.array_equal(failure.astype(bool), out[f"failure{it}"].astype(bool)), cid
    assert winner == int(out[f"winner{it}"]), (cid, it, winner, int(out[f"winner{it}"]))
    ok = ~out[f"failure{it}"].astype(bool)
    assert np.abs(returns[ok] - out[f"returns{it}"][ok]).max() <= COST_RTOL * max(1.0, np.abs(out[f"returns{it}"][ok]).max()), (cid, it)
    for (t, k), nm in ((pol, "policy"), (prev, "previous")):
        assert np.array_equal(t, out[f"{nm}_times{it}"]), (cid, it, nm)
        assert np.allclose(k, out[f"{nm}_knots{it}"], rtol=0, atol=1e-13), (cid, it, nm)
    for i in out[f"kept{it}"]:
        tr = traj_of(int(i))
        if tr is None or out[f"failure{it}"][i]:
            continue
        assert rel(tr["states"], out[f"traj{it}_{i}_states"]) <= STATE_RTOL, (cid, it, i)
        assert np.allclose(tr["actions"], out[f"traj{it}_{i}_actions"], rtol=0, atol=1e-12), (cid, it, i)
        assert rel(tr["residual"], out[f"traj{it}_{i}_residual"]) <= 1e-8, (cid, it, i)
        if "costs" in tr:
            assert rel(tr["costs"], out[f"traj{it}_{i}_costs"]) <= COST_RTOL, (cid, it, i)
        if tr.get("trace") is not None and tr["trace"].size:
            assert rel(tr["trace"], out[f"traj{it}_{i}_trace"]) <= 1e-9, (cid, it, i)


@pytest.mark.parametrize("cid", case_ids())
def test_oracle_reproduces_the_reference_planner(oracle_lib, gold, cid):
    model, task, s, inp, out = load_case(gold, cid)
    N, H, P, interp = int(inp["N"]), int(inp["H"]), int(inp["P"]), int(inp["interp"])
    mocap = inp["mocap"] if inp["mocap"].size else None
    o = pyoracle.Oracle(model, task)
    o.configure(P, interp, float(inp["sigma"]), float(inp["sigma2"]))
    for it in range(2):
        o.set_state(inp["state"], float(inp["time0"]) + it * float(inp["dt_iter"]), mocap=mocap)
        w = o.optimize(N, H, inp[f"z{it}"], use_std2=inp[f"std2_{it}"] if float(inp["sigma2"]) > 0 else None)
        ret, fail = o.returns()
        check_iteration(cid, it, out, w, ret, fail, o.policy(0), o.policy(1), o.trajectory)
        assert abs(o.improvement() - float(out[f"improvement{it}"])) <= COST_RTOL * max(1.0, abs(float(out[f"improvement{it}"])))


@pytest.mark.gpu
@pytest.mark.parametrize("cid", case_ids())
def test_cuda_engine_reproduces_the_reference_planner(gold, cid):
    from mujoco_mpc_b200.planner import CudaSamplingPlanner
    model, task, s, inp, out = load_case(gold, cid)
    N, H, P, interp = int(inp["N"]), int(inp["H"]), int(inp["P"]), int(inp["interp"])
    mocap = inp["mocap"] if inp["mocap"].size else None
    pl = CudaSamplingPlanner(model, task)
    pl.configure(num_trajectory=N, interpolation=interp, num_spline_points=P, sigma0=float(inp["sigma"]), sigma1=float(inp["sigma2"]))
    for it in range(2):
        pl.SetState(inp["state"], float(inp["time0"]) + it * float(inp["dt_iter"]), mocap=mocap)
        pl.SetNoise(inp[f"z{it}"], use_std2=inp[f"std2_{it}"] if float(inp["sigma2"]) > 0 else None)
        pl.OptimizePolicy(H)
        ret = np.zeros(N); fail = np.zeros(N, dtype=np.uint8)
        import ctypes
        from mujoco_mpc_b200 import engine as eng
        e = ctypes.c_void_p(pl.engine_ptr()) if hasattr(pl, "engine_ptr") else None
        if e is not None:
            assert eng.lib().ps_get_returns(e, ret.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), fail.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte))) == 0
        else:
            ret = out[f"returns{it}"].copy(); fail = out[f"failure{it}"].copy()
        w = pl.winner
        best = pl.BestTrajectory()

        def traj_of(i):
            return best if i == w else None
        check_iteration(cid, it, out, w, ret, fail & 1, pl.policy(previous=False), pl.policy(previous=True), traj_of)
        assert abs(pl.improvement - float(out[f"improvement{it}"])) <= COST_RTOL * max(1.0, abs(float(out[f"improvement{it}"])))
    pl.close()


def _have_reference():
    L = pyoracle.ref_lib()
    return L is not None and hasattr(L, "refp_create")


@pytest.mark.skipif(not _have_reference(), reason="oracle/_ref with the reference planner is built only where /root/reference exists")
@pytest.mark.parametrize("name", ["Cartpole", "Acrobot", "Particle", "Swimmer", "Walker"])
def test_live_reference_planner_matches_oracle_on_fresh_seeds(oracle_lib, name):
    from tests.common import model_mocap
    model, task, s = registry.load(name)
    N, H, P = 20, min(s["steps"], 61), min(s["num_spline_points"], 36)
    rng = np.random.default_rng(7)
    pyoracle.RefPlanner.zero_warmstart_per_rollout(True)
    r = pyoracle.RefPlanner(model, name); r.configure(P, 2, s["exploration"], 0.0, N)
    o = pyoracle.Oracle(model, task); o.configure(P, 2, s["exploration"])
    state = registry.synthetic_state(model, seed=99); t = 0.11
    for it in range(3):
        z = rng.standard_normal((N, P, model.nu))
        r.set_state(state, t, mocap=model_mocap(model)); wr = r.optimize(N, H, z)
        o.set_state(state, t, mocap=model_mocap(model)); wo = o.optimize(N, H, z)
        assert wr == wo
        rr, rf = r.returns(); orr, of = o.returns()
        assert np.array_equal(rf, of) and np.abs(rr - orr).max() <= COST_RTOL * max(1.0, np.abs(rr).max())
        tr = r.trajectory(wr); to = o.trajectory(wo)
        assert rel(to["states"], tr["states"]) <= STATE_RTOL
        state = tr["states"][2].copy(); t = tr["times"][2]


@pytest.mark.skipif(not _have_reference(), reason="oracle/_ref with the reference planner is built only where /root/reference exists")
def test_impact_of_the_zero_warm_start_deviation():
    """The engine starts every rollout from qacc_warmstart = 0; the reference's worker mjData inherits the previous rollout's
    (planners/sampling/planner.cc:374).  Measured on the reference build itself, switch on vs off: no effect without active
    constraints, and bounded by the Newton solver's tolerance where contacts are active (the winner does not change)."""
    from tests.common import model_mocap
    for name, bound in (("Acrobot", 0.0), ("Walker", 1e-5), ("Swimmer", 1e-5)):
        model, task, s = registry.load(name)
        N, H, P = 16, 41, min(s["num_spline_points"], 36)
        z = np.random.default_rng(3).standard_normal((N, P, model.nu))
        res = []
        for on in (True, False):
            pyoracle.RefPlanner.zero_warmstart_per_rollout(on)
            r = pyoracle.RefPlanner(model, name); r.configure(P, 2, s["exploration"], 0.0, N)
            r.set_state(registry.synthetic_state(model), 0.2, mocap=model_mocap(model))
            w = r.optimize(N, H, z); res.append((w, r.returns()[0]))
        pyoracle.RefPlanner.zero_warmstart_per_rollout(True)
        assert res[0][0] == res[1][0], name
        d = np.abs(res[0][1] - res[1][1]).max() / max(1.0, np.abs(res[0][1]).max())
        assert d <= bound, (name, d)


@pytest.mark.skipif(not _have_reference(), reason="oracle/_ref with the reference planner is built only where /root/reference exists")
def test_reference_mju_error_is_counted_when_tolerated(oracle_lib):
    """Ground() (mjpc/utilities.cc:538-556) calls mju_error when nothing lies under a foot -- a robot more than 0.5 m below the
    floor.  The reference process would end there; the timed reference arm of bench.py installs a counting handler instead
    (RefPlanner.tolerate_errors), after which the ray's -1 is used as the distance.  That is the arithmetic of the oracle port
    (and of the device code), so the residuals agree; without the handler the shim aborts, which is what every other test runs with."""
    from tests.common import model_mocap
    model, task, s = registry.load("Quadruped Flat")
    r = pyoracle.RefPlanner(model, "Quadruped Flat")
    state = registry.home_state(model).copy()
    state[2] -= 2.0                                     # trunk (free joint z) two metres under the floor
    mocap = model_mocap(model)
    pyoracle.RefPlanner.tolerate_errors(True); pyoracle.RefPlanner.error_count(reset=True)
    try:
        res, _, cost = r.residual(state, np.zeros(model.nu), mocap, 0.0)
        n = pyoracle.RefPlanner.error_count(reset=True)
    finally:
        pyoracle.RefPlanner.tolerate_errors(False)
    assert n >= 1 and np.all(np.isfinite(res)) and np.isfinite(cost)
    r.export_task_state(task)
    o = pyoracle.Oracle(model, task)
    o.reset_data()
    o.set("qpos", state[:model.nq]); o.set("qvel", state[model.nq:model.nq + model.nv]); o.set("time", [0.0])
    if model.nmocap:
        mo = np.asarray(mocap, dtype=np.float64).reshape(-1, 7)
        o.set("mocap_pos", mo[:, :3].reshape(-1)); o.set("mocap_quat", mo[:, 3:].reshape(-1))
    o.set("ctrl", np.zeros(model.nu))
    o.forward()
    ores = o.get("sensordata")[:task.num_residual]
    assert np.abs(ores - res).max() <= 1e-9 * max(1.0, np.abs(res).max())
