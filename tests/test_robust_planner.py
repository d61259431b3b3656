"""RobustPlanner (reference mjpc/planners/robust/robust_planner.cc:91-157 over a
SamplingPlanner delegate, Trajectory::NoisyRollout mjpc/trajectory.cc:100-210).

CPU: (1) the DEVICE code path of the force-perturbed rollout (csrc/ps_rollout.cuh + ps_smooth.cuh compiled by
tests/hostsim) against the oracle's noisy rollout under the same injected Ornstein-Uhlenbeck driving noise, bit for
bit; (2) the oracle's candidate scoring / selection against a numpy statement of robust_planner.cc:132-156.
GPU: CudaRobustPlanner against the oracle (written after this round's GPU budget was spent: see the marker)."""
import numpy as np
import pytest
from mujoco_mpc_b200 import registry, task as T
from oracle import pyoracle
from tests.common import model_mocap, run_isolated, synthetic_state
from tests.hostsim import driver


@pytest.fixture(scope="module")
def hostsim_lib(oracle_lib):
    return driver.build()


def setup(name, N, H, ncand, reps, std, rate, seed=3):
    model, task, s = registry.load(name)
    if name.startswith("Quadruped"):
        T.quadruped_snapshot(model, task, mode="Quadruped", gait="Trot", time=0.2)
    P, interp = min(s["num_spline_points"], 36), 2
    rng = np.random.default_rng(seed)
    z = rng.standard_normal((N, P, model.nu))
    w = rng.standard_normal((ncand * reps, H - 1, 6 * model.nbody))
    state = synthetic_state(model)
    o = pyoracle.Oracle(model, task); o.configure(P, interp, s["exploration"])
    shift = max((H - 1) * s["timestep"] / (P - 1), 1e-5)
    times = np.array([k * shift for k in range(P)])
    half = 0.5 * (model.actuator_ctrlrange[:, 1] - model.actuator_ctrlrange[:, 0])
    knots = 0.2 * half * rng.standard_normal((P, model.nu))
    o.set_state(state, 0.0, mocap=model_mocap(model)); o.set_policy(times, knots, interp)
    out = o.robust_optimize(N, H, z, ncand, reps, w, std, rate)
    return model, task, s, o, out, state, z, w, (times, knots), interp


@pytest.mark.parametrize("name,H,std", [("Cartpole", 41, 0.5), ("Walker", 31, 2.0), ("Quadruped Flat", 26, 1.0)])
def test_noisy_rollout_device_code_matches_oracle(hostsim_lib, name, H, std):
    N, ncand, reps, rate = 12, 3, 2, 0.1
    model, task, s, o, out, state, z, w, nominal, interp = setup(name, N, H, ncand, reps, std, rate)
    ret, _ = o.returns()
    for c in range(ncand):
        kt, kn = o.policy(2 + int(out["order"][c]))
        for j in range(reps):
            h = driver.rollout(model, task, state, 0.0, kt, kn, interp, H, s["timestep"], s["integrator"], xfrc=(w[c * reps + j], std, rate))
            assert bool(h["flags"] & 1) == bool(out["noisy_failure"][c, j])
            if not out["noisy_failure"][c, j]:
                assert h["total"] == out["noisy_returns"][c, j], (name, c, j, h["total"], out["noisy_returns"][c, j])
        # the perturbation matters: a noisy return differs from the candidate's clean return
        assert not np.all(out["noisy_returns"][c] == ret[out["order"][c]])
    # and with the perturbation switched off the device path reproduces the clean rollout
    kt, kn = o.policy(2 + int(out["order"][0]))
    h = driver.rollout(model, task, state, 0.0, kt, kn, interp, H, s["timestep"], s["integrator"])
    assert h["total"] == ret[out["order"][0]]


@pytest.mark.parametrize("name,H,std", [("Walker", 31, 2.0), ("Quadruped Flat", 26, 1.0), ("Cartpole", 41, 0.5)])
def test_noisy_rollout_phase_parallel_coding_matches_oracle(hostsim_lib, name, H, std):
    # the coding coop_robust_kernel runs (one lane here): per-dof accumulation of J' x instead of per-body Jacobians
    N, ncand, reps, rate = 12, 3, 2, 0.1
    model, task, s, o, out, state, z, w, nominal, interp = setup(name, N, H, ncand, reps, std, rate)
    ret, _ = o.returns()
    for c in range(ncand):
        kt, kn = o.policy(2 + int(out["order"][c]))
        for j in range(reps):
            h = driver.rollout(model, task, state, 0.0, kt, kn, interp, H, s["timestep"], s["integrator"], coop=True,
                               xfrc=(w[c * reps + j], std, rate))
            assert bool(h["flags"] & 1) == bool(out["noisy_failure"][c, j])
            if not out["noisy_failure"][c, j]:
                assert h["total"] == out["noisy_returns"][c, j], (name, c, j, h["total"], out["noisy_returns"][c, j])
    kt, kn = o.policy(2 + int(out["order"][0]))
    h = driver.rollout(model, task, state, 0.0, kt, kn, interp, H, s["timestep"], s["integrator"], coop=True)
    assert h["total"] == ret[out["order"][0]]


def test_robust_selection_rule(oracle_lib):
    N, H, ncand, reps, std, rate = 16, 31, 4, 3, 1.5, 0.1
    model, task, s, o, out, state, z, w, nominal, interp = setup("Walker", N, H, ncand, reps, std, rate, seed=8)
    ret, _ = o.returns()
    order = np.array(sorted(range(N), key=lambda i: (ret[i], i))[:ncand])
    assert np.array_equal(out["order"], order)
    scores = []
    for c in range(ncand):                    # robust_planner.cc:136-149: running mean seeded with the candidate's own score
        mean, valid = ret[order[c]], 0
        for j in range(reps):
            if out["noisy_failure"][c, j]:
                continue
            mean = (valid * mean + out["noisy_returns"][c, j]) / (valid + 1); valid += 1
        scores.append(mean)
    assert np.array_equal(out["scores"], np.array(scores))
    best = int(np.argmin(scores))             # first minimum, as `mean_return < best_score`
    assert out["best"] == best
    pt, pk = o.policy(0); ct, ck = o.policy(2 + int(order[best]))
    assert np.array_equal(pk, ck) and np.array_equal(pt, ct)
    # OptimizePolicyCandidates does not re-time the nominal (robust_planner.cc:91-95): knot times are untouched
    assert np.array_equal(pt, nominal[0])


def test_zero_force_noise_repeats_the_clean_return(oracle_lib):
    N, H, ncand, reps = 8, 21, 2, 2
    model, task, s, o, out, *_ = setup("Cartpole", N, H, ncand, reps, 0.0, 0.1)
    ret, _ = o.returns()
    for c in range(ncand):
        assert np.all(out["noisy_returns"][c] == ret[out["order"][c]])


# ---- GPU: the host class over the engine (green on hardware since the round-1 driver run)


@pytest.mark.gpu
@pytest.mark.parametrize("name,H,std,flavour,kernel", [
    ("Cartpole", 41, 0.5, 0, "tiny"), ("Walker", 31, 2.0, 0, "planar_coop"), ("Quadruped Flat", 26, 1.0, 0, "legged_coop"),
    ("Walker", 31, 2.0, 3, "planar"), ("Quadruped Flat", 26, 1.0, 3, "legged")])
def test_cuda_robust_planner_matches_oracle(oracle_lib, name, H, std, flavour, kernel):
    run_isolated("tests.test_robust_planner", "check_cuda_robust_planner_matches_oracle", name, H, std, flavour, kernel, timeout=240)


def check_cuda_robust_planner_matches_oracle(name, H, std, flavour, kernel):
    from mujoco_mpc_b200.planner import CudaRobustPlanner
    N, ncand, reps, rate = 64, 6, 4, 0.1
    model, task, s, o, out, state, z, w, nominal, interp = setup(name, N, H, ncand, reps, std, rate, seed=12)
    P = len(nominal[0])
    pl = CudaRobustPlanner(model, task, flavour=flavour)
    pl.configure(ncandidates=ncand, nrepetitions=reps, xfrc_std=std, xfrc_rate=rate)
    pl.delegate.configure(num_trajectory=N, interpolation=interp, num_spline_points=P, sigma0=s["exploration"], sigma1=0.0)
    pl.delegate.set_policy(nominal[0], nominal[1], interp)
    pl.delegate.SetNoise(z)
    pl.SetForceNoise(w, H)
    pl.SetState(state, 0.0, mocap=model_mocap(model))
    pl.OptimizePolicy(H)
    r = pl.results()
    ret, _ = o.returns()
    # the delegate's ranking and clean scores
    for c in range(ncand):
        assert abs(pl.delegate.CandidateScore(c) - ret[out["order"][c]]) <= 1e-8 * max(1.0, abs(ret[out["order"][c]]))
    assert np.array_equal((r["noisy_failure"] & 1).astype(bool), out["noisy_failure"].astype(bool))
    ok = ~out["noisy_failure"].astype(bool)
    assert np.allclose(r["noisy_returns"][ok], out["noisy_returns"][ok], rtol=1e-8, atol=1e-8)
    assert np.allclose(r["scores"], out["scores"], rtol=1e-8, atol=1e-8)
    assert pl.best_candidate == out["best"]
    pt, pk = pl.policy(); ot, ok_ = o.policy(0)
    assert np.array_equal(pt, ot) and np.allclose(pk, ok_, rtol=0, atol=1e-12)
    assert pl.delegate.winner == out["order"][out["best"]]
    assert pl.reroll_kernel() == kernel, pl.reroll_kernel()
    pl.close()


@pytest.mark.gpu
def test_cuda_robust_planner_device_noise_and_settings(oracle_lib):
    run_isolated("tests.test_robust_planner", "check_cuda_robust_planner_device_noise_and_settings", timeout=240)


def check_cuda_robust_planner_device_noise_and_settings():
    from mujoco_mpc_b200.planner import CudaRobustPlanner
    model, task, s = registry.load("Cartpole")
    pl = CudaRobustPlanner(model, task)
    st = pl.settings()
    # robust_planner.cc:48-57: defaults when the model has no robust_* numerics
    assert st["nrepetitions"] == 5 and st["xfrc_std"] == 0.1 and st["xfrc_rate"] == 0.1
    assert st["ncandidates"] == s["num_trajectory"] // 5
    pl.configure(ncandidates=8, nrepetitions=3, xfrc_std=1.0)
    pl.delegate.configure(num_trajectory=256)
    pl.SetSeed(5)
    pl.Reset(s["steps"], np.zeros(model.nu))
    state = synthetic_state(model)
    H = 51
    costs = []
    for k in range(4):
        pl.SetState(state, 0.0)
        pl.OptimizePolicy(H)
        r = pl.results()
        assert r["scores"].shape == (8,) and r["noisy_returns"].shape == (8, 3)
        assert 0 <= pl.best_candidate < 8 and pl.best_candidate == int(np.argmin(r["scores"]))
        # repetitions of one candidate see different forces
        assert len(np.unique(r["noisy_returns"][0])) == 3
        costs.append(r["scores"].min())
    # one candidate: no re-rolls, the delegate's winner is taken
    pl.configure(ncandidates=1)
    pl.SetState(state, 0.0); pl.OptimizePolicy(H)
    assert pl.best_candidate == 0 and pl.results()["scores"].size == 0
    a = pl.ActionFromPolicy(0.0)
    assert np.all(a >= model.actuator_ctrlrange[:, 0]) and np.all(a <= model.actuator_ctrlrange[:, 1])
    pl.close()
