"""Planner-level GPU tests through the drop-in host class (CudaSamplingPlanner): several
OptimizePolicy iterations in closed loop against the oracle's planner restatement under the same
injected noise, plus the reference's own planner-level assertions re-expressed
(sampling_planner_test.cc:58-107, agent_test.cc:204-262)."""
import numpy as np
import pytest
from mujoco_mpc_b200 import registry
from mujoco_mpc_b200.planner import CudaSamplingPlanner
from oracle import pyoracle
from tests.common import synthetic_state

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,interp", [("Cartpole", 0), ("Particle", 2), ("Acrobot", 1), ("Walker", 2)])
def test_closed_loop_iterations_match_oracle(oracle_lib, name, interp):
    model, task, s = registry.load(name)
    N, H, P, sigma = 96, min(s["steps"], 81), min(s["num_spline_points"], 36), s["exploration"]
    rng = np.random.default_rng(11)
    pl = CudaSamplingPlanner(model, task); pl.configure(num_trajectory=N, interpolation=interp, sigma0=sigma)
    o = pyoracle.Oracle(model, task); o.configure(P, interp, sigma)
    state = synthetic_state(model); t = 0.0
    for it in range(6):
        z = rng.standard_normal((N, P, model.nu))
        pl.SetState(state, t); pl.SetNoise(z); pl.OptimizePolicy(H)
        o.set_state(state, t); w = o.optimize(N, H, z)
        assert pl.winner == w, (it, pl.winner, w)
        for prev in (False, True):
            gt, gk = pl.policy(previous=prev); ot, ok = o.policy(1 if prev else 0)
            assert np.array_equal(gt, ot) and np.allclose(gk, ok, rtol=0, atol=1e-13), (it, prev)
        assert abs(pl.improvement - o.improvement()) <= 1e-8 * max(1.0, abs(o.improvement()))
        for tq in (t, t + 0.013, t + 5.0):
            assert np.allclose(pl.ActionFromPolicy(tq), o.action(tq), rtol=0, atol=1e-13)
            assert np.allclose(pl.ActionFromPolicy(tq, use_previous=True), o.action(tq, use_previous=True), rtol=0, atol=1e-13)
        bt = pl.BestTrajectory(); ot = o.trajectory(w)
        assert np.abs(bt["states"] - ot["states"]).max() <= 1e-9 * max(1.0, np.abs(ot["states"]).max())
        assert abs(bt["total_return"] - ot["total_return"]) <= 1e-8 * max(1.0, abs(ot["total_return"]))
        # advance the "plant" along the winner's trajectory by 3 steps
        state = ot["states"][3].copy(); t = ot["times"][3]
    pl.close()


def load_particle_fixture(timestep=None):
    import os
    from mujoco_mpc_b200 import mjcf, task as T
    m = mjcf.load(os.path.join(os.path.dirname(__file__), "fixtures", "particle_task.xml"))
    if timestep:
        m.v["opt_timestep"] = timestep; m._c = None
    return m, T.make_task(m, "ParticleTest")


def test_random_search_reaches_goal():
    # sampling_planner_test.cc:44-113, same fixture and settings: particle_task.xml, dt 0.1, 26 steps,
    # 10 samples, noise_exploration[0] = 0.01, 1000 OptimizePolicy iterations from the same state;
    # final best-trajectory state within 1e-1 of (goal, 0 velocity); actions inside ctrlrange
    model, task = load_particle_fixture(timestep=0.1)
    pl = CudaSamplingPlanner(model, task)
    pl.configure(sigma0=0.01)
    pl.SetSeed(3)
    steps = int(max(min(2.5 / 0.1 + 1, 512), 1))
    mocap = np.array([0.25, 0, 0.01, 1, 0, 0, 0.0])
    pl.SetState(np.zeros(4), 0.0, mocap=mocap)
    for _ in range(1000):
        pl.OptimizePolicy(steps)
    bt = pl.BestTrajectory()
    final = bt["states"][steps - 1]
    assert abs(final[0] - 0.25) < 1e-1 and abs(final[1] - 0.0) < 1e-1 and abs(final[2]) < 1e-1 and abs(final[3]) < 1e-1
    assert np.all(bt["actions"][:steps - 1] <= 1.0) and np.all(bt["actions"][:steps - 1] >= -1.0)
    pl.close()


def test_rollout_residual_state_alignment():
    # rollout_test.cc:127-144: for the copy-residual task, residual[t] == states[t] (minus the goal in
    # the position block) -- pins that the residual is evaluated at the PRE-step state
    model, task = load_particle_fixture()
    pl = CudaSamplingPlanner(model, task)
    pl.configure(num_trajectory=4, sigma0=0.3)
    pl.SetSeed(1)
    st = np.array([0.05, -0.02, 0.3, -0.1]); mocap = np.array([0.1, 0.2, 0.01, 1, 0, 0, 0.0])
    pl.SetState(st, 0.0, mocap=mocap); pl.OptimizePolicy(100)
    bt = pl.BestTrajectory()
    expect = bt["states"].copy(); expect[:, 0] -= 0.1; expect[:, 1] -= 0.2
    assert np.abs(expect - bt["residual"]).sum() < 1e-5
    pl.close()


def test_previous_policy_is_bit_exact():
    # agent_test.cc:204-262: ActionFromPolicy(use_previous=true) returns the pre-iteration policy exactly
    model, task = load_particle_fixture(timestep=0.1)
    pl = CudaSamplingPlanner(model, task); pl.configure(num_trajectory=128)
    pl.SetSeed(5)
    pl.SetState(np.zeros(4), 0.0); pl.OptimizePolicy(51)
    before = [pl.ActionFromPolicy(t) for t in (0.0, 0.1, 0.2)]
    pl.SetState(np.zeros(4), 0.0); pl.OptimizePolicy(51)
    after_prev = [pl.ActionFromPolicy(t, use_previous=True) for t in (0.0, 0.1, 0.2)]
    # previous_policy is the RESAMPLED nominal of the new iteration, which interpolates the old policy
    # at the same time origin: identical knot times => identical actions
    for a, b in zip(before, after_prev):
        assert np.array_equal(a, b)
    pl.close()


def test_ranked_candidates(oracle_lib):
    model, task, s = registry.load("Cartpole")
    N, H, P = 64, 51, 10
    z = np.random.default_rng(2).standard_normal((N, P, 1))
    pl = CudaSamplingPlanner(model, task); pl.configure(num_trajectory=N, interpolation=2, sigma0=0.5)
    state = synthetic_state(model)
    pl.SetState(state, 0.0); pl.SetNoise(z); pl.OptimizePolicy(H)        # establishes a nominal
    pl.SetState(state, 0.0); pl.SetNoise(z)
    n = pl.OptimizePolicyCandidates(5, H)
    assert n == 5
    scores = [pl.CandidateScore(c) for c in range(n)]
    assert scores == sorted(scores)
    a0 = pl.ActionFromCandidatePolicy(0, 0.2)
    pl.CopyCandidateToPolicy(0)
    assert np.array_equal(pl.ActionFromPolicy(0.2), a0)
    pl.close()
