"""CPU-side check of the DEVICE code path: the engine's kernel-source headers (csrc/ps_*.cuh) are
compiled with g++ (tests/hostsim) and one rollout per case is diffed against the oracle.  With
FMA contraction off on both sides and the shared deterministic sin/cos the two restatements agree
bit for bit, so the assertion is equality; the GPU run of the same source is covered by
tests/test_gpu_parity.py."""
import numpy as np
import pytest
from mujoco_mpc_b200 import registry, task as T
from oracle import pyoracle
from tests.hostsim import driver


@pytest.fixture(scope="module")
def hostsim_lib(oracle_lib):
    return driver.build()


def check(name, H, prep=None, N=3, sigma=0.2, time=0.3, seed=0, coop=False, spec=None, P=None, interp=None):
    m, t, s = registry.load(name)
    if prep:
        prep(m, t)
    P = min(s["num_spline_points"], 36) if P is None else P
    interp = s["interpolation"] if interp is None else interp
    st = registry.synthetic_state(m, seed=3)
    z = np.random.default_rng(seed).standard_normal((N, P * m.nu))
    o = pyoracle.Oracle(m, t); o.configure(P, interp, sigma); o.set_state(st, time)
    o.optimize(N, H, z)
    nt, nk = o.policy(1)
    for i in range(N):
        ot = o.trajectory(i)
        _, kn = o.policy(2 + i)
        if spec:
            h = driver.rollout_spec(m, t, st, time, nt, kn, interp, H, s["timestep"])
            assert h["variant"] == spec
        else:
            h = driver.rollout(m, t, st, time, nt, kn, interp, H, s["timestep"], s["integrator"], coop=coop)
        assert h["flags"] == (1 if ot["failure"] else 0)
        if ot["failure"]:
            continue
        assert np.array_equal(h["states"], ot["states"]), (name, i, np.abs(h["states"] - ot["states"]).max())
        assert np.array_equal(h["residual"], ot["residual"]), (name, i, np.abs(h["residual"] - ot["residual"]).max())
        assert h["total"] == ot["total_return"]


@pytest.mark.parametrize("name,H", [("Particle", 51), ("Cartpole", 101), ("Acrobot", 101), ("Swimmer", 40), ("Walker", 60)])
def test_kernel_source_matches_oracle(hostsim_lib, name, H):
    check(name, H)


# The edges of the domain: the terminal forward pass alone (H = 1), one integrated step, kMaxTrajectoryHorizon = 512
# (mjpc/trajectory.h:27), one knot, the 36-knot maximum of the reference's GUI (planners/sampling/planner.h:36), on every capacity
# class and both codings of the step.
@pytest.mark.parametrize("name,H,P,interp,coop", [
    ("Cartpole", 1, None, None, False), ("Cartpole", 2, None, None, False), ("Cartpole", 512, None, None, False),
    ("Particle", 51, 1, 0, False), ("Particle", 51, 2, 1, False), ("Acrobot", 40, 36, 2, False),
    ("Walker", 1, None, None, False), ("Walker", 2, None, None, True), ("Walker", 30, 36, 2, True),
    ("Quadruped Flat", 1, None, None, True), ("Quadruped Flat", 2, None, None, False), ("Quadruped Flat", 6, 36, 1, True),
    ("Quadruped Flat", 6, 1, 0, True)])
def test_kernel_source_matches_oracle_at_the_size_limits(hostsim_lib, name, H, P, interp, coop):
    prep = (lambda m, t: T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)) if name.startswith("Quadruped") else None
    check(name, H, prep=prep, P=P, interp=interp, coop=coop)


@pytest.mark.parametrize("name,H,sigma", [("Particle", 51, 0.2), ("Cartpole", 101, 0.2), ("Cartpole", 201, 1.0), ("Acrobot", 101, 0.2),
                                          ("Particle", 101, 1.0)])
def test_model_specialised_source_matches_oracle(hostsim_lib, name, H, sigma):
    # csrc/ps_variant_tiny_spec.cu's instantiations (compile-time tables, bounded tree walks, split row storage);
    # sigma = 1 drives the cart / the particle into their joint limits, i.e. through the constraint rows
    check(name, H, N=4, sigma=sigma, spec="tiny_" + name)


def test_capacity_overflow_is_rolled_again_at_large_capacity(hostsim_lib):
    """A quadruped lying on its side has 28 contacts / 102 rows -- beyond the scoring kernels' 24 contacts.  The engine's
    fix-up pass re-rolls such samples at the LeggedXLT capacities (64 / 256); that instantiation must equal the oracle
    (which has no capacity at all) bit for bit, like every other one."""
    m, t, s = registry.load("Quadruped Flat")
    T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    st = fallen_quadruped_state(m)
    P, interp, H = min(s["num_spline_points"], 36), 2, 31
    rng = np.random.default_rng(4)
    for trial in range(2):
        kn = rng.normal(0, 0.3, (P, m.nu)); nt = np.linspace(0.0, 0.3, P)
        o = pyoracle.Oracle(m, t); o.configure(P, interp, 0.0); o.set_state(st, 0.0); o.set_policy(nt, kn, interp)
        o.optimize(1, H, np.zeros((1, P, m.nu)))
        ot = o.trajectory(0)
        _, knots_used = o.policy(2)
        h = driver.rollout(m, t, st, 0.0, o.policy(1)[0], knots_used, interp, H, s["timestep"], s["integrator"], coop=True)
        assert "legged_xl" in h["variant"], h["variant"]
        assert h["flags"] == 0
        assert np.array_equal(h["states"], ot["states"]) and np.array_equal(h["residual"], ot["residual"]) and h["total"] == ot["total_return"]


def fallen_quadruped_state(m):
    st = registry.home_state(m)
    st[2] = 0.12; st[3:7] = [np.cos(np.pi / 4), np.sin(np.pi / 4), 0.0, 0.0]       # on its side, just above the floor
    return st


def test_model_specialisation_rejects_other_models(hostsim_lib):
    m, t, s = registry.load("Cartpole")
    m.v["jnt_limited"] = np.array([1, 1], dtype=np.int32); m._c = None
    with pytest.raises(RuntimeError, match="differs from the Cartpole specialisation"):
        driver.rollout_spec(m, t, registry.synthetic_state(m), 0.0, np.zeros(1), np.zeros((1, m.nu)), 0, 5, s["timestep"])
    m, t, s = registry.load("Walker")
    with pytest.raises(RuntimeError, match="sizes differ"):
        driver.rollout_spec(m, t, registry.synthetic_state(m), 0.0, np.zeros(1), np.zeros((1, m.nu)), 0, 5, s["timestep"])


def test_generated_spec_header_is_current():
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert subprocess.run([sys.executable, os.path.join(root, "tools", "gen_spec_header.py"), "--check"]).returncode == 0, \
        "csrc/ps_spec_generated.h is stale: run python tools/gen_spec_header.py"


@pytest.mark.parametrize("mode,gait,angvel", [("Quadruped", "Trot", 0.0), ("Biped", "Trot", 0.0), ("Walk", "Canter", 0.5),
                                               ("Scramble", "Gallop", 0.0), ("Flip", "Stand", 0.0)])
def test_kernel_source_matches_oracle_quadruped(hostsim_lib, mode, gait, angvel):
    check("Quadruped Flat", 60, prep=lambda m, t: T.quadruped_snapshot(
        m, t, mode=mode, gait=gait, time=0.3, position=(0.1, 0.2), heading=(0.5, 0.1), speed=0.5, angvel=angvel))


@pytest.mark.parametrize("mode,gait,angvel", [("Quadruped", "Trot", 0.0), ("Biped", "Trot", 0.0), ("Walk", "Canter", 0.5),
                                               ("Scramble", "Gallop", 0.0), ("Flip", "Stand", 0.0)])
def test_cooperative_source_matches_oracle_quadruped(hostsim_lib, mode, gait, angvel):
    # the warp-per-rollout engine (ps_coop*.cuh) compiled with a lane count of 1
    check("Quadruped Flat", 60, coop=True, prep=lambda m, t: T.quadruped_snapshot(
        m, t, mode=mode, gait=gait, time=0.3, position=(0.1, 0.2), heading=(0.5, 0.1), speed=0.5, angvel=angvel))


def test_cooperative_source_matches_oracle_walker(hostsim_lib):
    check("Walker", 80, coop=True)


def test_ground_ray_sees_props(oracle_lib):
    # Ground() (utilities.cc:538-556) must hit the box / ramp / hill scenery, not only the floor:
    # put the robot over the mocap box (top face at z = 0.3) and compare the Gait residual target.
    m, t, s = registry.load("Quadruped Flat")
    T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    st = registry.home_state(m)
    st[0], st[2] = -2.5, 0.56                      # above the box centre
    o = pyoracle.Oracle(m, t); o.configure(3, 2, 0.0); o.set_state(st, 0.1)
    o.optimize(1, 2, np.zeros((1, 3 * m.nu)))
    r_box = o.trajectory(0)["residual"][0]
    st2 = st.copy(); st2[0] = 0.0
    o.set_state(st2, 0.1); o.optimize(1, 2, np.zeros((1, 3 * m.nu)))
    r_floor = o.trajectory(0)["residual"][0]
    gait = slice(7, 11)
    moving = r_floor[gait] != 0
    assert moving.any()
    # same pose 0.3 + 0.01 higher ground => foot-height residual lower by the box height above the floor plane
    assert np.allclose(r_floor[gait][moving] - r_box[gait][moving], 0.31, atol=1e-12)
    h = driver.rollout(m, t, st, 0.1, np.array([0.1, 0.2, 0.3]), np.zeros((3, m.nu)), 2, 2, s["timestep"], s["integrator"])
    assert np.array_equal(h["residual"][0], r_box)
