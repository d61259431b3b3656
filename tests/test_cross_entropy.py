"""Cross-entropy planner (reference mjpc/planners/cross_entropy/planner.cc).

CPU: the oracle's CE restatement against an independent numpy statement of the update rule
(:196-278) on the candidates / returns the oracle itself produced.  GPU: CudaCrossEntropyPlanner
closed loop against the oracle under the same injected noise -- elite order identical, policy
(elite mean), variance, returns within the north_star tolerances."""
import numpy as np
import pytest
from mujoco_mpc_b200 import registry
from oracle import pyoracle
from tests.common import synthetic_state


def ce_update_numpy(knots, returns, n_elite):
    """cross_entropy/planner.cc:196-278 on candidate knots [N][P*nu] and returns [N]"""
    key = np.where(np.isnan(returns), 1e300, returns)
    order = np.array(sorted(range(len(returns)), key=lambda i: (key[i], i)), dtype=np.int64)
    mean = np.zeros(knots.shape[1]); avg = 0.0
    for e in range(n_elite):                              # sequential sums, elite order
        mean = mean + knots[order[e]]; avg += returns[order[e]]
    mean = mean * (1.0 / n_elite); avg /= n_elite
    diff = knots[order[0]] - mean                         # the reference measures the BEST candidate only (:246)
    var = np.zeros_like(mean)
    for _ in range(n_elite):
        var = var + diff * diff / (n_elite - 1)
    return order, mean, var, avg, max(avg - returns[order[0]], 0.0)


def run_oracle(model, task, s, N, n_elite, H, P, interp, z, state, t, policy=None, variance=None, std_min=0.05):
    o = pyoracle.Oracle(model, task)
    o.configure(P, interp, s["exploration"])
    o.ce_configure(std_min, n_elite, variance)
    o.set_state(state, t)
    if policy is not None:
        o.set_policy(policy[0], policy[1], interp)
    out = o.ce_optimize(N, H, z)
    return o, out


@pytest.mark.parametrize("name", ["Particle", "Cartpole"])
def test_oracle_update_rule(oracle_lib, name):
    model, task, s = registry.load(name)
    N, n_elite, H, P, interp = 24, 5, 31, min(s["num_spline_points"], 36), 2
    z = np.random.default_rng(5).standard_normal((N, P, model.nu))
    times = np.linspace(0, 0.3, P); knots = 0.1 * np.random.default_rng(6).standard_normal((P, model.nu))
    o, out = run_oracle(model, task, s, N, n_elite, H, P, interp, z, synthetic_state(model), 0.0, policy=(times, knots))
    cand = np.stack([o.policy(2 + i)[1].reshape(-1) for i in range(N)])
    ret, _ = o.returns()
    order, mean, var, avg, imp = ce_update_numpy(cand, ret, n_elite)
    assert np.array_equal(out["order"][:n_elite], order[:n_elite])
    _, pk = o.policy(0)
    assert np.array_equal(pk.reshape(-1), mean)
    assert np.array_equal(out["variance"].reshape(-1)[:mean.size], var)
    assert out["avg_return"] == avg and out["improvement"] == imp
    # candidates stay inside the control range and EVERY candidate is perturbed (:395-409)
    lo, hi = model.actuator_ctrlrange[:, 0], model.actuator_ctrlrange[:, 1]
    c3 = cand.reshape(N, P, model.nu)
    assert (c3 >= lo - 1e-15).all() and (c3 <= hi + 1e-15).all()
    _, nominal = o.policy(2)          # candidate 0 differs from the resampled nominal
    assert not np.array_equal(c3[0], c3[1])


def test_oracle_variance_feeds_next_iteration(oracle_lib):
    # noise std of iteration k+1 is max(sqrt(variance_k), std_min) (:360-363): with z = 1 everywhere
    # and no clamping, candidate - nominal equals that std
    model, task, s = registry.load("Particle")
    N, n_elite, H, P = 8, 3, 11, 4
    var_in = np.array([[1e-6, 4e-4], [9e-4, 1e-8], [2.5e-3, 1e-4], [0.0, 1e-2]])
    z = np.ones((N, P, model.nu))
    o, out = run_oracle(model, task, s, N, n_elite, H, P, 1, z, synthetic_state(model), 0.0,
                        policy=(np.linspace(0, 0.1, P), np.zeros((P, model.nu))), variance=var_in, std_min=0.01)
    _, c0 = o.policy(2)
    assert np.allclose(c0, np.maximum(np.sqrt(var_in), 0.01), rtol=0, atol=1e-15)


@pytest.mark.gpu
@pytest.mark.parametrize("name,interp,N,n_elite", [("Particle", 2, 64, 6), ("Cartpole", 1, 96, 9), ("Walker", 2, 64, 8)])
def test_gpu_cross_entropy_matches_oracle(oracle_lib, name, interp, N, n_elite):
    from mujoco_mpc_b200.planner import CudaCrossEntropyPlanner
    model, task, s = registry.load(name)
    H, P = min(s["steps"], 61), min(s["num_spline_points"], 36)
    std_min = 0.5 * s["exploration"]
    rng = np.random.default_rng(21)
    pl = CudaCrossEntropyPlanner(model, task)
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
        assert np.array_equal(gt, ot) and np.allclose(gk, ok, rtol=0, atol=1e-13), it
        assert np.allclose(pl.variance(P), out["variance"][:P], rtol=1e-12, atol=1e-300), it
        assert abs(pl.avg_return - out["avg_return"]) <= 1e-8 * max(1.0, abs(out["avg_return"]))
        assert abs(pl.improvement - out["improvement"]) <= 1e-8 * max(1.0, abs(out["improvement"]))
        for tq in (t, t + 0.017, t + 9.0):
            assert np.allclose(pl.ActionFromPolicy(tq), o.action(tq), rtol=0, atol=1e-13)
        bt = pl.BestTrajectory()      # the NOMINAL trajectory (:426-428)
        assert abs(bt["total_return"] - out["nominal_return"]) <= 1e-8 * max(1.0, abs(out["nominal_return"]))
        best = o.trajectory(int(out["order"][0]))
        state = best["states"][2].copy(); t = best["times"][2]
    pl.close()


@pytest.mark.gpu
def test_gpu_elite_ranking_large_batch(oracle_lib):
    # device ranking (K4) against numpy at a size the oracle cannot roll out: order, mean, variance
    from mujoco_mpc_b200.engine import Engine
    model, task, s = registry.load("Cartpole")
    N, n_elite, H, P = 20000, 700, 21, 6
    e = Engine(model, task)
    e.set_state(synthetic_state(model), 0.0)
    e.set_nominal(np.linspace(0, 0.2, P), np.zeros((P, model.nu)), 2)
    e.set_seed(7)
    e.set_param_std(np.full((P, model.nu), 0.3))
    e.rollouts(N + 1, H, 0.0)
    mean, var, avg, best, order = e.elite_update(N, n_elite)
    ret, _ = e.returns()
    cand = np.stack([e.knots(int(i)).reshape(-1) for i in order])
    ref_order = np.array(sorted(range(N), key=lambda i: (ret[i], i))[:n_elite])
    assert np.array_equal(order, ref_order)
    m = np.zeros(P * model.nu)
    for k in range(n_elite):
        m = m + cand[k]
    m = m * (1.0 / n_elite)
    assert np.array_equal(mean.reshape(-1), m)
    assert best == ret[ref_order[0]]
    assert np.array_equal(e.knots(N), np.zeros((P, model.nu)))       # the nominal column stays un-noised
    e.close()
