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
from tests.common import synthetic_state
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
    o.set_state(state, 0.0); o.set_policy(times, knots, interp)
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
