"""GPU parity: the CUDA engine (through the C ABI) against the CPU oracle under the SAME injected
knot noise.  Tolerances are the north_star's: per-step qpos/qvel 1e-9 relative, total trajectory
cost 1e-8 relative, identical best-sample index."""
import numpy as np
import pytest
from mujoco_mpc_b200 import registry
from mujoco_mpc_b200.engine import Engine
from tests.common import synthetic_state, oracle_iteration

pytestmark = pytest.mark.gpu

STATE_RTOL, COST_RTOL = 1e-9, 1e-8


def rel_err(a, b):
    return np.abs(a - b).max() / max(1.0, np.abs(b).max())


def run_case(name, N, interp, H=None, seed=2024, sigma=None, time=0.37, warm=True, prep=None, engine=None, expect=None, state=None, P=None):
    model, task, s = registry.load(name)
    if prep is not None:
        prep(model, task)
    H = H or s["steps"]
    P = min(s["num_spline_points"], 36) if P is None else P
    sigma = s["exploration"] if sigma is None else sigma
    rng = np.random.default_rng(seed)
    z = rng.standard_normal((N, P, model.nu))
    state = synthetic_state(model) if state is None else state(model)
    nominal = None
    if warm:   # steady state: a non-trivial previous plan, offset in time so resampling matters
        nt = time - 0.013 + np.linspace(0, (H - 1) * s["timestep"], P)
        nk = rng.uniform(-0.6, 0.6, (P, model.nu)) * 0.5 * (model.actuator_ctrlrange[:, 1] - model.actuator_ctrlrange[:, 0])
        nominal = (nt, nk)
    o, nt, nk, w = oracle_iteration(model, task, s, state, time, N, H, P, interp, sigma, z, nominal=nominal)
    import os
    if engine:                       # pin one kernel flavour (read by ps_create)
        os.environ["PS_ENGINE"] = engine
    try:
        e = Engine(model, task)
    finally:
        os.environ.pop("PS_ENGINE", None)
    e.set_state(state, time)
    e.set_nominal(nt, nk, interp)
    e.set_noise(z)
    gw, gwr, gnr = e.rollouts(N, H, sigma, writeback=True)
    if expect:
        assert e.variant() == expect, (e.variant(), expect)
    oret, ofail = o.returns()
    gret, gfail = e.returns()
    assert np.array_equal(ofail.astype(bool), (gfail & 1).astype(bool))
    # bit 1 = the rollout ran out of contact / constraint-row capacity.  The engine rolls such samples again at the large
    # capacities of its fix-up pass (coop_fixup_kernel), so no sample may still carry it: every sample is in the parity set
    # (a rollout that DIVERGED -- bit 0, scored 1e6 on both sides -- may still carry it: its contact set exploded on the way)
    capped = ((gfail & 2) != 0) & ((gfail & 1) == 0)
    assert not capped.any(), (name, int(capped.sum()))
    ok = ~ofail.astype(bool) & ~capped
    assert np.abs(gret[ok] - oret[ok]).max() <= COST_RTOL * max(1.0, np.abs(oret[ok]).max()), (name, np.abs(gret - oret).max())
    if not capped.any():
        assert gw == w, (name, gw, w, gret[gw], oret[w])
    else:
        assert gw == int(np.argmin(np.where(capped, gret, oret))), (name, gw, w)
    if not capped[gw]:
        assert abs(gwr - oret[gw]) <= COST_RTOL * max(1.0, abs(oret[gw]))
    assert abs(gnr - oret[0]) <= COST_RTOL * max(1.0, abs(oret[0]))
    # every step of every sample
    tr = e.all_trajectories()
    worst_s = worst_r = 0.0
    for i in range(N):
        ot = o.trajectory(i)
        if ot["failure"] or capped[i]:
            continue
        worst_s = max(worst_s, rel_err(tr["states"][i], ot["states"]))
        worst_r = max(worst_r, rel_err(tr["residual"][i], ot["residual"]))
        assert np.allclose(tr["actions"][i], ot["actions"], rtol=0, atol=1e-12)
    assert worst_s <= STATE_RTOL, (name, worst_s)
    assert worst_r <= 1e-8, (name, worst_r)
    # candidate knots and single-trajectory read-back
    pt, pk = o.policy(2 + gw)
    assert np.allclose(e.knots(gw), pk, rtol=0, atol=1e-14)
    e.close()
    return worst_s, worst_r


# The edges of the domain, through the C ABI: one sample, the terminal forward pass alone (H = 1), one integrated step, a batch
# that is not a multiple of a warp, kMaxTrajectoryHorizon = 512 (mjpc/trajectory.h:27), one knot, the 36-knot maximum of the
# reference's GUI (planners/sampling/planner.h:36) -- on every capacity class.  (tests/test_hostsim_parity.py runs the same edges
# on the host-compiled kernel source at bit equality.)
@pytest.mark.parametrize("name,N,H,interp,P,warm", [
    ("Cartpole", 1, 1, 0, None, False), ("Cartpole", 1, 2, 2, None, True), ("Cartpole", 33, 7, 1, None, True),
    ("Cartpole", 7, 512, 2, None, True), ("Particle", 129, 51, 0, 1, False), ("Acrobot", 5, 40, 2, 36, True),
    ("Walker", 3, 1, 2, None, False), ("Walker", 3, 2, 2, None, True), ("Quadruped Flat", 3, 1, 2, None, False),
    ("Quadruped Flat", 3, 2, 2, None, True), ("Quadruped Flat", 37, 6, 1, 36, True)])
def test_parity_at_the_size_limits(oracle_lib, name, N, H, interp, P, warm):
    from mujoco_mpc_b200 import task as T
    prep = (lambda m, t: T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)) if name.startswith("Quadruped") else None
    run_case(name, N, interp, H=H, P=P, warm=warm, prep=prep)


def test_sizes_beyond_the_limits_are_rejected():
    """kMaxTrajectoryHorizon = 512 steps and 36 knots are hard limits of the reference's arrays; the engine refuses anything
    beyond them (and empty batches / horizons) with an error instead of truncating."""
    from mujoco_mpc_b200.engine import EngineError
    model, task, s = registry.load("Cartpole")
    e = Engine(model, task)
    e.set_state(synthetic_state(model), 0.0)
    with pytest.raises(EngineError):
        e.set_nominal(np.linspace(0, 1, 37), np.zeros((37, model.nu)), 2)
    e.set_nominal(np.linspace(0, 1, 5), np.zeros((5, model.nu)), 2)
    for N, H in ((4, 513), (0, 10), (4, 0)):
        with pytest.raises(EngineError):
            e.rollouts(N, H, 0.1)
    e.rollouts(4, 512, 0.1)                   # the limit itself is fine
    e.close()


@pytest.mark.parametrize("name", ["Particle", "Cartpole", "Acrobot"])
@pytest.mark.parametrize("interp", [0, 1, 2])
def test_parity_small_models(oracle_lib, name, interp):
    run_case(name, N=256, interp=interp)


def test_parity_config2_sizes(oracle_lib):
    # BASELINE config 2: Acrobot + Particle, 4,096 samples, cubic
    run_case("Acrobot", N=4096, interp=2, H=101)
    run_case("Particle", N=4096, interp=2)


@pytest.mark.parametrize("name,N", [("Swimmer", 512), ("Walker", 1024)])
def test_parity_planar_models(oracle_lib, name, N):
    # Swimmer: implicit integrator + fluid forces + activation states + joint limits;
    # Walker: capsule-plane contacts, pyramidal cones, Newton solver
    run_case(name, N=N, interp=2)


QUAD_CASES = [  # mode, gait, walk angvel: every branch of quadruped.cc:33-221
    ("Quadruped", "Trot", 0.0), ("Quadruped", "Stand", 0.0), ("Biped", "Trot", 0.0), ("Walk", "Walk", 0.0),
    ("Walk", "Canter", 0.5), ("Scramble", "Gallop", 0.0), ("Flip", "Stand", 0.0)]


@pytest.mark.parametrize("mode,gait,angvel", QUAD_CASES)
def test_parity_quadruped(oracle_lib, mode, gait, angvel):
    # free joint, 12 torque actuators, elliptic cones with condim-6 feet, frictionloss + joint limits,
    # ground ray casts in the gait residual
    from mujoco_mpc_b200 import task as T

    def prep(model, task):
        T.quadruped_snapshot(model, task, mode=mode, gait=gait, time=0.3, position=(0.1, 0.2), heading=(0.5, 0.1),
                             speed=0.5, angvel=angvel)
    ws, wr = run_case("Quadruped Flat", N=192, interp=2, prep=prep)
    assert ws <= STATE_RTOL


def test_parity_quadruped_one_second_horizon(oracle_lib):
    # BASELINE config 4 geometry: 1 s horizon at the task's 10 ms agent timestep
    from mujoco_mpc_b200 import task as T
    run_case("Quadruped Flat", N=512, interp=2, H=101,
             prep=lambda m, t: T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.37))


@pytest.mark.parametrize("name,engine,expect", [
    ("Cartpole", "warp", "tiny_coop"), ("Cartpole", "thread", "tiny_Cartpole"), ("Cartpole", "dense", "tiny"), ("Acrobot", "warp", "tiny_coop"),
    ("Acrobot", "thread", "tiny_Acrobot"), ("Acrobot", "dense", "tiny"), ("Particle", "thread", "tiny_Particle"), ("Particle", "dense", "tiny"),
    ("Walker", "warp", "planar_coop"), ("Walker", "thread", "planar"),
    ("Quadruped Flat", "warp", "legged_coop"), ("Quadruped Flat", "thread", "legged_sparse"), ("Quadruped Flat", "dense", "legged")])
def test_parity_every_kernel_flavour(oracle_lib, name, engine, expect):
    # the engine picks a kernel by batch size; here each flavour is pinned and checked on the same batch
    from mujoco_mpc_b200 import task as T
    prep = (lambda m, t: T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.37)) if name.startswith("Quadruped") else None
    run_case(name, N=160, interp=2, H=41, prep=prep, engine=engine, expect=expect)


def test_model_specialised_kernel_falls_back_on_any_integer_difference(oracle_lib):
    # the model-specialised lean kernels (csrc/ps_variant_tiny_spec.cu) are compiled for the registered tasks' integers;
    # a cost specification or a model that differs in any of them must run on the generic lean kernel, same results
    def other_norm(m, t):
        t.norm[2] = 2; t.num_norm_parameter[2] = 1                      # velocity term: quadratic -> L2 with one parameter
        npar = sum(int(t.num_norm_parameter[k]) for k in range(2))
        for k in range(31, npar, -1):
            t.norm_parameter[k] = t.norm_parameter[k - 1]
        t.norm_parameter[npar] = 0.1
    run_case("Cartpole", N=96, interp=2, H=41, prep=other_norm, expect="tiny")

    def limited_pole(m, t):
        m.v["jnt_limited"] = np.array([1, 1], dtype=np.int32); m.v["jnt_range"] = np.array([[-1.8, 1.8], [-0.3, 0.3]]); m._c = None
    run_case("Cartpole", N=96, interp=2, H=41, prep=limited_pole, expect="tiny")
    run_case("Cartpole", N=96, interp=2, H=41, expect="tiny_Cartpole")


@pytest.mark.parametrize("engine,expect", [("warp", "legged_coop"), ("thread", "legged_sparse"), ("dense", "legged")])
def test_capacity_overflow_does_not_change_the_answer(oracle_lib, engine, expect):
    # a quadruped lying on its side: 28 contacts / 102 rows, beyond every scoring kernel's capacity (24 or 16 contacts).
    # The reference never truncates a rollout (trajectory.cc:169-173); the fix-up pass must reproduce the oracle -- which has
    # no capacity limit -- on every sample and pick the same winner
    from mujoco_mpc_b200 import task as T
    from tests.test_hostsim_parity import fallen_quadruped_state
    run_case("Quadruped Flat", N=96, interp=2, H=31, prep=lambda m, t: T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.37),
             engine=engine, expect=expect, state=fallen_quadruped_state)


def test_first_iteration_zero_nominal(oracle_lib):
    run_case("Cartpole", N=128, interp=0, warm=False)


def test_reroll_single_trajectory(oracle_lib):
    model, task, s = registry.load("Cartpole")
    N, H, P = 64, 51, 10
    z = np.random.default_rng(1).standard_normal((N, P, model.nu))
    state = synthetic_state(model)
    o, nt, nk, w = oracle_iteration(model, task, s, state, 0.0, N, H, P, 2, 0.5, z)
    e = Engine(model, task)
    e.set_state(state, 0.0); e.set_nominal(nt, nk, 2); e.set_noise(z)
    gw, _, _ = e.rollouts(N, H, 0.5)             # cost-only mode
    assert gw == w
    g = e.trajectory(w); ot = o.trajectory(w)     # re-rolled with writeback
    assert rel_err(g["states"], ot["states"]) <= STATE_RTOL
    assert rel_err(g["trace"], ot["trace"]) <= 1e-9
    assert abs(g["total_return"] - ot["total_return"]) <= COST_RTOL * max(1, abs(ot["total_return"]))
    assert np.allclose(g["times"], ot["times"], rtol=0, atol=1e-12)
    e.close()


def test_traces_of_all_samples(oracle_lib):
    # SamplingPlanner::Traces (planner.cc:388-434) draws trace[i][t] -> trace[i][t+1] for every sample:
    # the writeback batch exposes exactly those points
    model, task, s = registry.load("Walker")
    N, H, P = 48, 31, 3
    z = np.random.default_rng(3).standard_normal((N, P, model.nu))
    state = synthetic_state(model)
    o, nt, nk, w = oracle_iteration(model, task, s, state, 0.0, N, H, P, 2, 0.5, z)
    e = Engine(model, task)
    e.set_state(state, 0.0); e.set_nominal(nt, nk, 2); e.set_noise(z)
    e.rollouts(N, H, 0.5, writeback=True)
    tr = e.all_traces()
    assert tr.shape == (N, H, 3 * task.num_trace) and task.num_trace > 0
    for i in range(N):
        ot = o.trajectory(i)
        if not ot["failure"]:
            assert rel_err(tr[i], ot["trace"]) <= 1e-9
    e.close()


def test_device_rng_statistics():
    model, task, s = registry.load("Particle")
    e = Engine(model, task)
    N, P = 65536, 5
    e.set_state(registry.home_state(model), 0.0)
    e.set_nominal(np.linspace(0, 0.5, P), np.zeros((P, model.nu)), 2)
    e.set_seed(7)
    e.rollouts(N, 3, 0.1)
    k = np.stack([e.knots(i) for i in range(0, 4096)])
    assert np.all(k[0] == 0)                        # sample 0 is the un-noised nominal
    zs = k[1:].ravel() / 0.1                        # ctrlrange +-1 => scale 1
    assert abs(zs.mean()) < 0.02 and abs(zs.std() - 1) < 0.02
    e.set_seed(7); e.rollouts(N, 3, 0.1)
    assert np.array_equal(e.knots(17), k[17])       # counter-based: reproducible
    e.close()
