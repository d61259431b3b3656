"""Physical-law checks of the oracle's restated MuJoCo step (plan item 4).

MuJoCo 3.1.4 is not in this container and the reference tree holds no `mj_step` golden vectors, so
the restated physics cannot be diffed against the real library ("parity unpinned", DESIGN.md
section 4).  These tests pin it to what any correct rigid-body step must satisfy instead:
the gravity term of the bias force is the gradient of the potential energy, the equations of
motion close (M qacc = sum of generalized forces), the Newton solver's output is a stationary
point of its own cost, a conservative system keeps its energy, and a body in free flight falls
with g while internal forces leave its momentum alone."""
import numpy as np
import pytest
from mujoco_mpc_b200 import registry
from oracle import pyoracle


def fresh(name, seed=5, vel=True):
    m, t, s = registry.load(name)
    o = pyoracle.Oracle(m, t)
    st = registry.synthetic_state(m, seed=seed, pos=0.3, vel=0.5 if vel else 0.0)
    return m, t, s, o, st


def load_state(o, m, st, ctrl=None):
    o.reset_data()
    o.set("qpos", st[:m.nq]); o.set("qvel", st[m.nq:m.nq + m.nv]); o.set("time", [0.0])
    o.set("ctrl", np.zeros(m.nu) if ctrl is None else ctrl)


def potential(o, m):
    x = np.array(o.get("xipos")[:3 * m.nbody]).reshape(m.nbody, 3)
    return -float(np.sum(m.body_mass[:, None] * x * np.asarray(m.opt_gravity)[None, :]))


def field(o, name, n):
    return np.array(o.get(name)[:n])


@pytest.mark.parametrize("name", ["Acrobot", "Cartpole", "Walker", "Swimmer"])
def test_gravity_bias_is_potential_gradient(oracle_lib, name):
    # at zero velocity qfrc_bias = dV/dq (hinge / slide models: dq is a plain coordinate increment)
    m, t, s, o, st = fresh(name, vel=False)
    st[m.nq:] = 0
    load_state(o, m, st); o.forward()
    bias = field(o, "qfrc_bias", m.nv)
    eps = 1e-6
    grad = np.zeros(m.nv)
    for j in range(m.njnt):
        assert m.jnt_type[j] in (2, 3)
        qa, da = m.jnt_qposadr[j], m.jnt_dofadr[j]
        v = []
        for sgn in (+1, -1):
            s2 = st.copy(); s2[qa] += sgn * eps
            load_state(o, m, s2); o.forward(); v.append(potential(o, m))
        grad[da] = (v[0] - v[1]) / (2 * eps)
    scale = max(1.0, np.abs(grad).max())
    assert np.abs(bias - grad).max() <= 2e-7 * scale, (name, bias, grad)


@pytest.mark.parametrize("name", ["Particle", "Cartpole", "Acrobot", "Swimmer", "Walker", "Quadruped Flat"])
def test_equations_of_motion_close(oracle_lib, name):
    m, t, s, o, st = fresh(name)
    if name.startswith("Quadruped"):
        from mujoco_mpc_b200 import task as T
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
        o = pyoracle.Oracle(m, t)
    ctrl = 0.4 * (m.actuator_ctrlrange[:, 1] - m.actuator_ctrlrange[:, 0]) * np.random.default_rng(2).uniform(-0.5, 0.5, m.nu)
    load_state(o, m, st, ctrl)
    for _ in range(5):                     # a few steps in, so warm starts and contacts are live
        o.step()
    o.forward()
    nv = m.nv
    M = field(o, "qM", nv * nv).reshape(nv, nv)
    assert np.array_equal(M, M.T) and np.all(np.linalg.eigvalsh(M) > 0)
    smooth = field(o, "qfrc_smooth", nv)
    assert np.allclose(smooth, field(o, "qfrc_passive", nv) - field(o, "qfrc_bias", nv) + field(o, "qfrc_actuator", nv), rtol=0, atol=1e-12 * max(1, np.abs(smooth).max()))
    fscale = max(1.0, np.abs(smooth).max())
    assert np.abs(M @ field(o, "qacc_smooth", nv) - smooth).max() <= 1e-10 * fscale
    # constrained: M qacc = qfrc_smooth + J' f to the solver's tolerance (a stationary point of its cost)
    qacc = field(o, "qacc", nv); qc = field(o, "qfrc_constraint", nv)
    resid = M @ qacc - smooth - qc
    assert np.abs(resid).max() <= 1e-5 * fscale, (name, np.abs(resid).max(), fscale)
    nefc = int(o.get("nefc")[0])
    if nefc:
        J = field(o, "efc_J", nefc * nv).reshape(nefc, nv); f = field(o, "efc_force", nefc)
        assert np.abs(J.T @ f - qc).max() <= 1e-9 * max(1.0, np.abs(qc).max())


def test_acrobot_energy_is_conserved_without_damping(oracle_lib):
    m, t, s = registry.load("Acrobot")
    m.v["dof_damping"] = np.zeros(m.nv); m.v["opt_timestep"] = 1e-4; m._c = None
    o = pyoracle.Oracle(m, t)
    st = np.array([2.0, -1.0, 0.5, -0.3])
    load_state(o, m, st)

    def energy():
        o.forward()
        v = field(o, "qvel", m.nv); M = field(o, "qM", m.nv * m.nv).reshape(m.nv, m.nv)
        return 0.5 * v @ M @ v + potential(o, m)
    e0 = energy()
    for _ in range(5000):                  # 0.5 s of swinging
        o.step()
    e1 = energy()
    assert abs(e1 - e0) <= 2e-3 * max(1.0, abs(e0)), (e0, e1)     # semi-implicit Euler: O(dt) drift


def test_quadruped_free_flight_momentum(oracle_lib):
    # lifted 2 m above the floor: no contacts; joint damping / friction loss / actuators are internal forces
    from mujoco_mpc_b200 import task as T
    m, t, s = registry.load("Quadruped Flat")
    T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    m.v["opt_timestep"] = 1e-3; m._c = None
    o = pyoracle.Oracle(m, t)
    st = registry.synthetic_state(m, seed=9, pos=0.05, vel=1.0)
    st[2] += 2.0
    ctrl = 0.3 * m.actuator_ctrlrange[:, 1] * np.random.default_rng(3).uniform(-1, 1, m.nu)
    load_state(o, m, st, ctrl)
    root = int(np.argmax(np.asarray(m.body_dofnum) == 6))     # the trunk (bodies 1, 2 are mocap props)
    dt = 1e-3

    def com():
        o.forward()
        return field(o, "subtree_com", 3 * m.nbody).reshape(m.nbody, 3)[root].copy()
    pts = [com()]
    assert int(o.get("ncon")[0]) == 0
    n = 200
    for _ in range(n):
        o.step(); pts.append(com())
    assert int(o.get("ncon")[0]) == 0
    pts = np.array(pts)
    # the centre of mass of the whole robot follows a parabola whatever the joints do:
    # second differences equal g dt^2 (to the integrator's O(dt) error), first differences change linearly
    acc = (pts[2:] - 2 * pts[1:-1] + pts[:-2]) / dt ** 2
    g = np.asarray(m.opt_gravity)
    assert np.abs(acc.mean(axis=0) - g).max() <= 2e-2, acc.mean(axis=0)
    assert np.abs(acc - g[None, :]).max() <= 0.5, np.abs(acc - g[None, :]).max()
    v_first = (pts[1] - pts[0]) / dt; v_last = (pts[-1] - pts[-2]) / dt
    assert np.abs(v_last - (v_first + g * (n - 1) * dt)).max() <= 5e-3
