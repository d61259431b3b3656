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


# ---- closed-form known answers of MuJoCo's soft-constraint model (solref = [timeconst, dampratio], solimp = [dmin,
# dmax, width, midpoint, power]): stiffness k = 1 / (dmax^2 timeconst^2 dampratio^2), impedance d(r) on the published
# sigmoid, regularisation R = (1 - d) / d * diagApprox, reference acceleration aref = -b v - k d r, force
# f = (aref - J qacc) / R.  These pin the constraint formulas of the restatement to numbers derived on paper.
def _compile(xml):
    from mujoco_mpc_b200 import mjcf, task as T
    m = mjcf.compile_xml(xml)
    return m, T.make_task(m, "Particle")


_SENSORS = """
  <sensor>
    <user name="Position" dim="3" user="0 1.0 0.0 1.0"/>
    <framepos name="position" objtype="site" objname="tip"/>
    <framelinvel name="velocity" objtype="site" objname="tip"/>
  </sensor>"""


def _impedance(r, dmin=0.9, dmax=0.95, width=0.001, mid=0.5, power=2.0):
    x = min(abs(r) / width, 1.0)
    a = 1 / mid ** (power - 1); b = 1 / (1 - mid) ** (power - 1)
    y = a * x ** power if x <= mid else 1 - b * (1 - x) ** power
    return dmin + y * (dmax - dmin)


def test_resting_sphere_sinks_by_the_closed_form_depth(oracle_lib):
    # frictionless sphere (1 kg, free joint) at rest on a plane with the default solver parameters: m g = f gives
    # |r| = g (1 - d(r)) / (d(r)^2 k), the fixed point below -- 0.367 mm, MuJoCo's familiar default sinking depth
    m, t = _compile("""
<mujoco model="rest">
  <option timestep="0.002"/>
  <worldbody>
    <geom name="floor" type="plane" size="1 1 .1" condim="1"/>
    <body name="ball" pos="0 0 0.1">
      <freejoint/>
      <geom name="ball" type="sphere" size="0.1" mass="1" condim="1"/>
      <site name="tip" pos="0 0 0"/>
    </body>
  </worldbody>""" + _SENSORS + "</mujoco>")
    o = pyoracle.Oracle(m, t)
    o.reset_data(); o.set("qpos", np.array(m.qpos0, dtype=float)); o.set("qvel", np.zeros(6)); o.set("time", [0.0])
    for _ in range(3000):
        o.step()
    depth = 0.1 - o.get("qpos")[2]
    k = 1.0 / (0.95 ** 2 * 0.02 ** 2)
    r = 4e-4
    for _ in range(200):
        d = _impedance(r); r = 9.81 * (1 - d) / (d * d * k)
    assert abs(r - 3.6718e-4) < 1e-8                                    # the number itself
    assert abs(depth - r) < 1e-12, (depth, r)
    assert int(o.get("ncon")[0]) == 1 and abs(o.get("efc_force")[0] - 9.81) < 1e-9
    assert np.abs(np.array(o.get("qvel")[:6])).max() < 1e-12


def test_friction_loss_first_step_acceleration(oracle_lib):
    # a 2 kg slider with frictionloss 3 N pushed by gravity along its axis, at rest.  Inside the friction bound the row
    # is quadratic with aref = 0 and R = (1 - dmin) / dmin / m, so qacc = (1 - dmin) F / m; beyond it the force
    # saturates at the loss: qacc = (F - loss) / m.
    def qacc(g):
        m, t = _compile(f"""
<mujoco model="loss">
  <option timestep="0.002" gravity="{g} 0 0"/>
  <worldbody>
    <body name="block" pos="0 0 0.5">
      <joint name="x" type="slide" axis="1 0 0" frictionloss="3"/>
      <geom name="block" type="box" size="0.1 0.1 0.1" mass="2" contype="0" conaffinity="0"/>
      <site name="tip" pos="0 0 0"/>
    </body>
  </worldbody>""" + _SENSORS + "</mujoco>")
        o = pyoracle.Oracle(m, t)
        o.reset_data(); o.set("qpos", [0.0]); o.set("qvel", [0.0]); o.set("time", [0.0]); o.forward()
        return o.get("qacc")[0], o.get("efc_force")[0]
    a, f = qacc(1.0)                       # F = 2 N < 3 N
    assert abs(a - 0.1 * 2.0 / 2.0) < 1e-9 and abs(f + 0.9 * 2.0) < 1e-9
    a, f = qacc(5.0)                       # F = 10 N > 3 N
    assert abs(a - (10.0 - 3.0) / 2.0) < 1e-9 and abs(f + 3.0) < 1e-9


def test_joint_limit_equilibrium_depth(oracle_lib):
    # a 0.5 kg slider resting against its lower limit under gravity: same fixed point with diagApprox = 1 / m
    m, t = _compile("""
<mujoco model="limit">
  <option timestep="0.002"/>
  <worldbody>
    <body name="block" pos="0 0 0.5">
      <joint name="z" type="slide" axis="0 0 1" limited="true" range="-0.05 0.2" damping="0.5"/>
      <geom name="block" type="box" size="0.1 0.1 0.1" mass="0.5" contype="0" conaffinity="0"/>
      <site name="tip" pos="0 0 0"/>
    </body>
  </worldbody>""" + _SENSORS + "</mujoco>")
    o = pyoracle.Oracle(m, t)
    o.reset_data(); o.set("qpos", [0.0]); o.set("qvel", [0.0]); o.set("time", [0.0])
    for _ in range(4000):
        o.step()
    depth = -0.05 - o.get("qpos")[0]
    k = 1.0 / (0.95 ** 2 * 0.02 ** 2)
    r = 4e-4
    for _ in range(200):
        d = _impedance(r); r = 9.81 * (1 - d) / (d * d * k)
    assert abs(depth - r) < 1e-11, (depth, r)
    assert abs(o.get("efc_force")[0] - 0.5 * 9.81) < 1e-8


def test_euler_implicit_damping_recursion_and_armature(oracle_lib):
    # mj_Euler with implicit joint damping: v' = v + h (F - b v) / (m + armature + h b), q' = q + h v'
    mass, arm, b, g, h = 2.0, 0.3, 4.0, 3.0, 0.01
    m, t = _compile(f"""
<mujoco model="damp">
  <option timestep="{h}" gravity="{g} 0 0"/>
  <worldbody>
    <body name="block" pos="0 0 0.5">
      <joint name="x" type="slide" axis="1 0 0" damping="{b}" armature="{arm}"/>
      <geom name="block" type="box" size="0.1 0.1 0.1" mass="{mass}" contype="0" conaffinity="0"/>
      <site name="tip" pos="0 0 0"/>
    </body>
  </worldbody>""" + _SENSORS + "</mujoco>")
    o = pyoracle.Oracle(m, t)
    o.reset_data(); o.set("qpos", [0.0]); o.set("qvel", [0.5]); o.set("time", [0.0])
    q, v = 0.0, 0.5
    for _ in range(200):
        o.step()
        v = v + h * (mass * g - b * v) / (mass + arm + h * b); q = q + h * v
    assert abs(o.get("qvel")[0] - v) < 1e-13 and abs(o.get("qpos")[0] - q) < 1e-13
    assert abs(v - mass * g / b) < 0.05                       # close to the terminal velocity F / b after 2 s


def test_filter_activation_dynamics(oracle_lib):
    # dyntype filter: act_dot = (ctrl - act) / tau, act' = act + h act_dot; force = gain * act on a general actuator
    tau, gain, h = 0.25, 3.0, 0.01
    m, t = _compile(f"""
<mujoco model="act">
  <option timestep="{h}" gravity="0 0 0"/>
  <worldbody>
    <body name="block" pos="0 0 0.5">
      <joint name="x" type="slide" axis="1 0 0"/>
      <geom name="block" type="box" size="0.1 0.1 0.1" mass="1" contype="0" conaffinity="0"/>
      <site name="tip" pos="0 0 0"/>
    </body>
  </worldbody>
  <actuator>
    <general name="a" joint="x" gainprm="{gain}" dyntype="filter" dynprm="{tau}" ctrllimited="true" ctrlrange="-1 1"/>
  </actuator>""" + _SENSORS + "</mujoco>")
    o = pyoracle.Oracle(m, t)
    o.reset_data(); o.set("qpos", [0.0]); o.set("qvel", [0.0]); o.set("act", [0.0]); o.set("time", [0.0]); o.set("ctrl", [0.8])
    act, v = 0.0, 0.0
    for _ in range(100):
        o.step()
        v = v + h * gain * act                                # force uses the activation BEFORE it is advanced
        act = act + h * (0.8 - act) / tau
    assert abs(o.get("act")[0] - act) < 1e-14 and abs(o.get("qvel")[0] - v) < 1e-13
    assert abs(act - 0.8 * (1 - (1 - h / tau) ** 100)) < 1e-12


def test_viscous_drag_of_the_inertia_box(oracle_lib):
    # inertia-box fluid model, viscosity only: force = -3 pi D mu v, torque = -pi D^3 mu w, D = mean box dimension; for a
    # solid box of half sizes (a, b, c) the equivalent inertia box is the box itself
    import math
    a, b, c, mu = 0.1, 0.2, 0.3, 0.7
    m, t = _compile(f"""
<mujoco model="drag">
  <option timestep="0.002" gravity="0 0 0" viscosity="{mu}"/>
  <worldbody>
    <body name="block" pos="0 0 0.5">
      <freejoint/>
      <geom name="block" type="box" size="{a} {b} {c}" mass="1.5" contype="0" conaffinity="0"/>
      <site name="tip" pos="0 0 0"/>
    </body>
  </worldbody>""" + _SENSORS + "</mujoco>")
    o = pyoracle.Oracle(m, t)
    vel = np.array([0.4, -0.3, 0.2, 0.5, 0.1, -0.6])       # free joint: linear (world), angular (local); body frame = world here
    o.reset_data(); o.set("qpos", np.array(m.qpos0, dtype=float)); o.set("qvel", vel); o.set("time", [0.0]); o.forward()
    D = (2 * a + 2 * b + 2 * c) / 3
    f = np.array(o.get("qfrc_passive")[:6])
    assert np.allclose(f[:3], -3 * math.pi * D * mu * vel[:3], rtol=1e-12, atol=1e-14)
    assert np.allclose(f[3:], -math.pi * D ** 3 * mu * vel[3:], rtol=1e-12, atol=1e-14)
