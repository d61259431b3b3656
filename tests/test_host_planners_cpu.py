"""The product's host planner classes (mujoco_mpc_b200/host/*.cc: CudaSamplingPlanner, CudaRobustPlanner),
compiled unchanged and linked against tests/hostsim/hostengine.cc -- a host-compiled stand-in for the engine
ABI whose rollouts are the engine's own kernel-source headers run through g++ -- against the oracle.

This is what can be said on a machine without a GPU about the host control flow: knot times and resampling,
candidate ranking, the layout of what is handed to ps_rollout_candidates, the running-mean selection and which
trajectory BestTrajectory() returns.  The CUDA launches themselves are covered by the `-m gpu` tests."""
import ctypes
import numpy as np
import pytest
from mujoco_mpc_b200 import planner, registry, task as T
from oracle import pyoracle
from tests.common import model_mocap, synthetic_state
from tests.hostsim import driver


@pytest.fixture()
def hostlib(oracle_lib, monkeypatch):
    L = planner.prototype(ctypes.CDLL(driver.build_planner_host()))
    monkeypatch.setattr(planner, "_LIB", L)
    return L


def case(name, N, H, seed):
    model, task, s = registry.load(name)
    if name.startswith("Quadruped"):
        T.quadruped_snapshot(model, task, mode="Quadruped", gait="Trot", time=0.2)
    P, interp = min(s["num_spline_points"], 36), 2
    rng = np.random.default_rng(seed)
    z = rng.standard_normal((N, P, model.nu))
    shift = max((H - 1) * s["timestep"] / (P - 1), 1e-5)
    times = np.array([k * shift for k in range(P)])
    half = 0.5 * (model.actuator_ctrlrange[:, 1] - model.actuator_ctrlrange[:, 0])
    knots = 0.2 * half * rng.standard_normal((P, model.nu))
    return model, task, s, P, interp, z, times, knots, synthetic_state(model), rng


@pytest.mark.parametrize("name,H,sliding,interp", [("Cartpole", 31, 0, 2), ("Walker", 21, 0, 2), ("Cartpole", 31, 1, 2), ("Walker", 21, 1, 1),
                                                    ("Particle", 21, 1, 0), ("Particle", 21, 0, 0)])
def test_sampling_planner_iterations_match_oracle(hostlib, name, H, sliding, interp):
    N = 10
    model, task, s, P, _, z, times, knots, state, rng = case(name, N, H, seed=21)
    mocap = model_mocap(model)
    o = pyoracle.Oracle(model, task); o.configure(P, interp, s["exploration"]); o.set_sliding_plan(sliding)
    pl = planner.CudaSamplingPlanner(model, task)
    pl.configure(num_trajectory=N, interpolation=interp, num_spline_points=P, sigma0=s["exploration"], sigma1=0.0, sliding=sliding)
    pl.Reset(H, np.zeros(model.nu)); o.set_policy(np.zeros(1), np.zeros((1, model.nu)), interp)      # Reset(): one knot at t = 0
    t = 0.0
    first = []
    for it in range(6):                       # planning iterations: resampling / sliding of the previous winner included
        zi = rng.standard_normal((N, P, model.nu))
        o.set_state(state, t, mocap=mocap); ow = o.optimize(N, H, zi)
        pl.SetNoise(zi); pl.SetState(state, t, mocap=mocap); pl.OptimizePolicy(H)
        ret, _ = o.returns()
        assert pl.winner == ow
        ot, ok = o.policy(0); pt, pk = pl.policy()
        assert np.array_equal(ot, pt) and np.array_equal(ok, pk), (name, it)
        assert len(pt) == P
        # a resampled plan starts at the planning time; a sliding plan keeps older nodes
        assert pt[0] <= t if sliding else pt[0] == t
        first.append(pt[0])
        assert pl.improvement == o.improvement()
        bt = pl.BestTrajectory()
        assert bt["total_return"] == ret[ow]
        for tq in (t, t + 0.37 * (H - 1) * s["timestep"]):
            assert np.array_equal(pl.ActionFromPolicy(tq), o.action(tq))
        t += (0.45 if it % 2 else 0.04) * (H - 1) * s["timestep"]      # the long hops drop nodes of a sliding plan
    if sliding and P > 6:
        assert first[-1] > first[0]           # nodes behind the planning time were discarded
    pl.close()


@pytest.mark.parametrize("name,H,std", [("Cartpole", 31, 0.5), ("Walker", 21, 2.0), ("Quadruped Flat", 16, 1.0)])
def test_robust_planner_matches_oracle(hostlib, name, H, std):
    N, ncand, reps, rate = 12, 4, 3, 0.1
    model, task, s, P, interp, z, times, knots, state, rng = case(name, N, H, seed=33)
    w = rng.standard_normal((ncand * reps, H - 1, 6 * model.nbody))
    o = pyoracle.Oracle(model, task); o.configure(P, interp, s["exploration"])
    mocap = model_mocap(model)
    o.set_state(state, 0.0, mocap=mocap); o.set_policy(times, knots, interp)
    out = o.robust_optimize(N, H, z, ncand, reps, w, std, rate)
    ret, _ = o.returns()

    pl = planner.CudaRobustPlanner(model, task)
    pl.configure(ncandidates=ncand, nrepetitions=reps, xfrc_std=std, xfrc_rate=rate)
    pl.delegate.configure(num_trajectory=N, interpolation=interp, num_spline_points=P, sigma0=s["exploration"], sigma1=0.0)
    pl.delegate.set_policy(times, knots, interp)
    pl.delegate.SetNoise(z)
    pl.SetForceNoise(w, H)
    pl.SetState(state, 0.0, mocap=mocap)
    pl.OptimizePolicy(H)
    r = pl.results()
    for c in range(ncand):
        assert pl.delegate.CandidateScore(c) == ret[out["order"][c]]
    assert np.array_equal((r["noisy_failure"] & 1).astype(bool), out["noisy_failure"].astype(bool))
    ok = ~out["noisy_failure"].astype(bool)
    assert np.array_equal(r["noisy_returns"][ok], out["noisy_returns"][ok])
    assert np.array_equal(r["scores"], out["scores"])
    assert pl.best_candidate == out["best"]
    assert pl.delegate.winner == out["order"][out["best"]]
    pt, pk = pl.policy(); ot, okn = o.policy(0)
    assert np.array_equal(pt, ot) and np.array_equal(pk, okn)
    assert np.array_equal(pt, times)                     # OptimizePolicyCandidates does not re-time the plan
    # BestTrajectory() is the delegate's clean rollout of the candidate that was picked
    bt = pl.BestTrajectory()
    assert bt["total_return"] == ret[out["order"][out["best"]]]
    pl.close()


def test_robust_planner_settings_and_single_candidate(hostlib):
    model, task, s = registry.load("Cartpole")
    pl = planner.CudaRobustPlanner(model, task)
    st = pl.settings()
    # robust_planner.cc:48-57: defaults when the model carries no robust_* numerics
    assert st["nrepetitions"] == 5 and st["xfrc_std"] == 0.1 and st["xfrc_rate"] == 0.1
    assert st["ncandidates"] == s["num_trajectory"] // 5
    N, H, P = 6, 21, 1
    pl.configure(ncandidates=1)
    pl.delegate.configure(num_trajectory=N)
    pl.Reset(H, np.zeros(model.nu))                   # one knot at t = 0: the plan RobustPlanner keeps (it never resamples)
    z = np.random.default_rng(2).standard_normal((N, P, model.nu))
    pl.delegate.SetNoise(z)
    pl.SetState(synthetic_state(model), 0.0)
    pl.OptimizePolicy(H)
    assert pl.best_candidate == 0 and pl.results()["scores"].size == 0
    t, k = pl.policy()
    assert len(t) == 1 and t[0] == 0.0
    a = pl.ActionFromPolicy(0.3)
    assert np.array_equal(a, np.clip(k[0], model.actuator_ctrlrange[:, 0], model.actuator_ctrlrange[:, 1]))
    pl.close()


@pytest.mark.parametrize("name,interp,N,n_elite", [("Particle", 2, 16, 4), ("Cartpole", 1, 20, 5), ("Walker", 2, 12, 3)])
def test_cross_entropy_planner_matches_oracle(hostlib, name, interp, N, n_elite):
    # the closed loop of tests/test_cross_entropy.py::test_gpu_cross_entropy_matches_oracle, on the host stand-in
    model, task, s = registry.load(name)
    H, P = min(s["steps"], 31), min(s["num_spline_points"], 36)
    std_min = 0.5 * s["exploration"]
    rng = np.random.default_rng(21)
    pl = planner.CudaCrossEntropyPlanner(model, task)
    pl.configure(num_trajectory=N, n_elite=n_elite, interpolation=interp, std_initial=s["exploration"], std_min=std_min)
    pl.Reset(H)
    o = pyoracle.Oracle(model, task); o.configure(P, interp, s["exploration"]); o.ce_configure(std_min, n_elite)
    times = np.linspace(0, (H - 1) * s["timestep"], P); knots = np.zeros((P, model.nu))
    pl.set_policy(times, knots, interp); o.set_policy(times, knots, interp)
    state = synthetic_state(model); t = 0.0
    for it in range(4):
        z = rng.standard_normal((N, P, model.nu))
        pl.SetState(state, t); pl.SetNoise(z); pl.OptimizePolicy(H)
        o.set_state(state, t); out = o.ce_optimize(N, H, z)
        assert np.array_equal(pl.order(n_elite), out["order"][:n_elite]), it
        gt, gk = pl.policy(0); ot, ok = o.policy(0)
        assert np.array_equal(gt, ot) and np.array_equal(gk, ok), it
        assert np.array_equal(pl.variance(P), out["variance"][:P]), it
        assert pl.avg_return == out["avg_return"] and pl.improvement == out["improvement"]
        for tq in (t, t + 0.017, t + 9.0):
            assert np.array_equal(pl.ActionFromPolicy(tq), o.action(tq))
        bt = pl.BestTrajectory()      # the NOMINAL trajectory (cross_entropy/planner.cc:426-428)
        assert bt["total_return"] == out["nominal_return"]
        best = o.trajectory(int(out["order"][0]))
        state = best["states"][2].copy(); t = best["times"][2]
    pl.close()


@pytest.mark.parametrize("task_name,interp,N,steps", [("Cartpole", 0, 10, 60), ("Acrobot", 2, 8, 24), ("Particle", 2, 8, 24),
                                                    ("Swimmer", 2, 6, 16), ("Walker", 2, 6, 16), ("Quadruped Flat", 2, 6, 40)])
def test_testspeed_closed_loop_matches_oracle(hostlib, monkeypatch, task_name, interp, N, steps):
    # mujoco_mpc_b200/testspeed.py (planner AND plant behind the C ABI) on the host stand-in, against the reference's
    # testspeed loop restated on the oracle (tests/common.py: oracle_closed_loop); the GPU run of the same loop is
    # tests/test_testspeed_closed_loop.py
    from mujoco_mpc_b200 import engine
    from mujoco_mpc_b200.testspeed import plant_model, synchronous_planning_cost
    from tests.common import oracle_closed_loop
    monkeypatch.setattr(engine, "_LIB", engine.prototype(hostlib))
    cache = {}

    def noise(k, n, P, nu):
        if k not in cache:
            cache[k] = np.random.default_rng(1000 + k).standard_normal((n, P, nu))
        return cache[k]
    total_time = steps * plant_model(task_name).opt_timestep
    g = synchronous_planning_cost(task_name, steps_per_planning_iteration=4, total_time=total_time, num_trajectory=N, interpolation=interp, noise=noise)
    o = oracle_closed_loop(task_name, 4, total_time, N, interp, noise)
    assert g["plans"] == o["plans"] == steps // 4
    assert g["time"] == o["time"]
    # exact on the contact tasks too: the plant hands the solver's warm start from one launch to the next
    # (ps_get_warmstart / ps_set_warmstart), as the reference's one mjData carries qacc_warmstart from step to step
    assert np.array_equal(g["state"], o["state"])
    assert abs(g["total_cost"] - o["total_cost"]) <= 1e-13 * max(1.0, abs(o["total_cost"]))      # summed in segments there


def test_reference_robust_planner_test_random_search(hostlib):
    # mjpc/test/planners/robust/robust_planner_test.cc:47-127 re-expressed: particle_task.xml, keyframe "ctrl_test",
    # Reset() with the keyframe's ctrl, dt 0.1, 26 steps, default robust settings (2 candidates x 5 repetitions,
    # xfrc 0.1), 1000 iterations from the same state; same expectations.  Noise is numpy's (the stand-in has no
    # generator); the reference draws from absl, so only the test's own tolerances are comparable.
    import os
    from mujoco_mpc_b200 import mjcf
    m = mjcf.load(os.path.join(os.path.dirname(__file__), "fixtures", "particle_task.xml"))
    m.v["opt_timestep"] = 0.1; m._c = None
    t = T.make_task(m, "ParticleTest")
    pl = planner.CudaRobustPlanner(m, t)
    st = pl.settings()
    assert (st["ncandidates"], st["nrepetitions"], st["xfrc_std"], st["xfrc_rate"]) == (2, 5, 0.1, 0.1)
    key = m.name2id("key", "ctrl_test")
    pl.Reset(512, m.key_ctrl[key])
    a = pl.ActionFromPolicy(2.0)                       # "some arbitrary, hard-coded time"
    assert abs(a[0] - 0.1) < 1e-4 and abs(a[1] - 0.2) < 1e-4
    steps = int(max(min(2.5 / 0.1 + 1, 512), 1))
    state = np.concatenate([m.key_qpos[key], m.key_qvel[key]])
    mocap = model_mocap(m)
    pl.SetState(state, 0.0, mocap=mocap)
    rng = np.random.default_rng(0)
    N, P = 10, 1                                       # sampling_trajectories default; the plan Reset() left: one node
    for _ in range(1000):
        pl.delegate.SetNoise(rng.standard_normal((N, P, m.nu)))
        pl.SetForceNoise(rng.standard_normal((2 * 5, steps - 1, 6 * m.nbody)), steps)
        pl.OptimizePolicy(steps)
    bt = pl.BestTrajectory()
    final = bt["states"][steps - 1]
    assert abs(final[0] - mocap[0]) < 1e-1 and abs(final[1] - mocap[1]) < 1e-1
    assert abs(final[2]) < 1e-1 and abs(final[3]) < 1e-1
    lo, hi = m.actuator_ctrlrange[:, 0], m.actuator_ctrlrange[:, 1]
    assert np.all(bt["actions"][:steps - 1] <= hi) and np.all(bt["actions"][:steps - 1] >= lo)
    pl.close()


def _particle_fixture(timestep=None):
    import os
    from mujoco_mpc_b200 import mjcf
    m = mjcf.load(os.path.join(os.path.dirname(__file__), "fixtures", "particle_task.xml"))
    if timestep:
        m.v["opt_timestep"] = timestep; m._c = None
    return m, T.make_task(m, "ParticleTest")


def test_reference_sampling_planner_test_random_search(hostlib):
    # mjpc/test/sampling_planner/sampling_planner_test.cc:44-113 re-expressed on the stand-in (numpy noise): dt 0.1,
    # 26 steps, 10 samples, exploration 0.01, 1000 iterations from the same state; the test's own expectations
    m, t = _particle_fixture(timestep=0.1)
    pl = planner.CudaSamplingPlanner(m, t)
    pl.configure(sigma0=0.01)
    steps = int(max(min(2.5 / 0.1 + 1, 512), 1))
    mocap = model_mocap(m)
    pl.SetState(np.zeros(4), 0.0, mocap=mocap)
    rng = np.random.default_rng(3)
    P = 11                                              # sampling_spline_points of the fixture
    for _ in range(1000):
        pl.SetNoise(rng.standard_normal((10, P, m.nu)))
        pl.OptimizePolicy(steps)
    bt = pl.BestTrajectory()
    final = bt["states"][steps - 1]
    assert abs(final[0] - mocap[0]) < 1e-1 and abs(final[1] - mocap[1]) < 1e-1 and abs(final[2]) < 1e-1 and abs(final[3]) < 1e-1
    assert np.all(bt["actions"][:steps - 1] <= 1.0) and np.all(bt["actions"][:steps - 1] >= -1.0)
    pl.close()


def test_reference_rollout_test_residual_state_alignment(hostlib):
    # mjpc/test/agent/rollout_test.cc:127-144: for the copy-residual task residual[t] == states[t] (minus the goal in the
    # position block): the residual is evaluated at the PRE-step state
    m, t = _particle_fixture()
    pl = planner.CudaSamplingPlanner(m, t)
    pl.configure(num_trajectory=4, sigma0=0.3)
    st = np.array([0.05, -0.02, 0.3, -0.1]); mocap = np.array([0.1, 0.2, 0.01, 1, 0, 0, 0.0])
    pl.SetNoise(np.random.default_rng(1).standard_normal((4, 11, m.nu)))
    pl.SetState(st, 0.0, mocap=mocap); pl.OptimizePolicy(100)
    bt = pl.BestTrajectory()
    expect = bt["states"].copy(); expect[:, 0] -= 0.1; expect[:, 1] -= 0.2
    assert np.abs(expect - bt["residual"]).sum() < 1e-5
    pl.close()


def test_reference_agent_test_previous_policy(hostlib):
    # mjpc/test/agent/agent_test.cc:204-262: ActionFromPolicy(use_previous=true) returns the pre-iteration policy exactly
    m, t = _particle_fixture(timestep=0.1)
    pl = planner.CudaSamplingPlanner(m, t); pl.configure(num_trajectory=16)
    rng = np.random.default_rng(5)
    pl.SetNoise(rng.standard_normal((16, 11, m.nu))); pl.SetState(np.zeros(4), 0.0); pl.OptimizePolicy(51)
    before = [pl.ActionFromPolicy(tq) for tq in (0.0, 0.1, 0.2)]
    pl.SetNoise(rng.standard_normal((16, 11, m.nu))); pl.SetState(np.zeros(4), 0.0); pl.OptimizePolicy(51)
    for a, tq in zip(before, (0.0, 0.1, 0.2)):
        assert np.array_equal(a, pl.ActionFromPolicy(tq, use_previous=True))
    pl.close()


def test_sampling_planner_second_exploration_std(hostlib):
    # noise_exploration[1] > 0: a sample uses the second std when its Bernoulli(0.2) draw says so (planner.cc:319-325);
    # the draw is injected as a per-sample flag on both sides
    N, H = 12, 31
    model, task, s, P, interp, z, times, knots, state, rng = case("Cartpole", N, H, seed=44)
    flags = (rng.uniform(size=N) < 0.4).astype(np.uint8); flags[0] = 0
    assert flags.any() and not flags.all()
    o = pyoracle.Oracle(model, task); o.configure(P, interp, 0.2, 0.9)
    o.set_policy(times, knots, interp); o.set_state(state, 0.0)
    w = o.optimize(N, H, z, use_std2=flags)
    pl = planner.CudaSamplingPlanner(model, task)
    pl.configure(num_trajectory=N, interpolation=interp, num_spline_points=P, sigma0=0.2, sigma1=0.9)
    pl.set_policy(times, knots, interp)
    pl.SetNoise(z, use_std2=flags); pl.SetState(state, 0.0); pl.OptimizePolicy(H)
    assert pl.winner == w
    ot, ok = o.policy(0); pt, pk = pl.policy()
    assert np.array_equal(ot, pt) and np.array_equal(ok, pk)
    # the flag matters: without it the candidates differ
    o2 = pyoracle.Oracle(model, task); o2.configure(P, interp, 0.2, 0.9)
    o2.set_policy(times, knots, interp); o2.set_state(state, 0.0); o2.optimize(N, H, z)
    assert not np.array_equal(o2.returns()[0], o.returns()[0])
    pl.close()
