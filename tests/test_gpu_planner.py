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


def test_random_search_reaches_goal():
    # sampling_planner_test.cc:44-113 on the in-tree particle model (fixed mocap goal variant),
    # with device-generated noise: final state within 1e-1 of the goal, actions within ctrlrange
    model, task, s = registry.load("Particle")
    from mujoco_mpc_b200 import task as T
    task.task_id = T.TASK_IDS["ParticleFixed"]
    pl = CudaSamplingPlanner(model, task); pl.configure(num_trajectory=128, interpolation=0, num_spline_points=11, sigma0=0.05)
    pl.SetSeed(3)
    state = np.zeros(4); goal = np.array([0.1, 0.1])
    mocap = np.array([goal[0], goal[1], 0.01, 1, 0, 0, 0.0])
    H = 26
    for _ in range(150):
        pl.SetState(state, 0.0, mocap=mocap); pl.OptimizePolicy(H)
    bt = pl.BestTrajectory()
    assert np.linalg.norm(bt["states"][-1][:2] - goal) < 1e-1
    assert np.all(bt["actions"] >= -1 - 1e-12) and np.all(bt["actions"] <= 1 + 1e-12)
    pl.close()


def test_previous_policy_is_bit_exact():
    # agent_test.cc:204-262: ActionFromPolicy(use_previous=true) returns the pre-iteration policy exactly
    model, task, s = registry.load("Particle")
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
