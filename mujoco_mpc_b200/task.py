"""PsTask: cost specification + residual snapshot crossing the C ABI (include/ps_engine.h).

Restates what ``mjpc::Task::Reset`` parses out of the model's leading ``user`` sensors and
``residual_*`` custom numerics (/root/reference/mjpc/task.cc:38-64,147-245) and what
``BaseResidualFn::Update`` snapshots each planning iteration (task.cc:112-123).
"""
import ctypes
import numpy as np
from . import psmodel as P

MAX_COST_TERMS, MAX_TASK_PARAMS, MAX_TASK_AUX = 128, 64, 128
TASK_IDS = {"Particle": 0, "ParticleFixed": 1, "ParticleTest": 2, "Cartpole": 3, "Acrobot": 4,
            "Swimmer": 5, "Walker": 6, "Quadruped": 7}
SPLINE_ZERO, SPLINE_LINEAR, SPLINE_CUBIC = 0, 1, 2
NORM_PARAM_DIM = {-1: 0, 0: 0, 1: 2, 2: 1, 3: 1, 5: 1, 6: 1, 7: 2, 8: 1}   # norm.cc:24-47


class PsTaskC(ctypes.Structure):
    _fields_ = [("task_id", ctypes.c_int), ("num_residual", ctypes.c_int), ("num_term", ctypes.c_int),
                ("num_trace", ctypes.c_int), ("num_parameters", ctypes.c_int),
                ("dim_norm_residual", ctypes.c_int * MAX_COST_TERMS),
                ("num_norm_parameter", ctypes.c_int * MAX_COST_TERMS),
                ("norm", ctypes.c_int * MAX_COST_TERMS),
                ("weight", ctypes.c_double * MAX_COST_TERMS),
                ("norm_parameter", ctypes.c_double * (2 * MAX_COST_TERMS)),
                ("risk", ctypes.c_double),
                ("parameters", ctypes.c_double * MAX_TASK_PARAMS),
                ("aux", ctypes.c_double * MAX_TASK_AUX),
                ("trace_sensor", ctypes.c_int * 4),
                ("sensor_id", ctypes.c_int * 16),
                ("iaux", ctypes.c_int * 16)]

# sensors each task's residual reads by name (SensorByName call sites in mjpc/tasks/<task>/<task>.cc)
TASK_SENSORS = {"Particle": ["position", "velocity"], "ParticleFixed": ["position", "velocity"],
                "Swimmer": ["nose", "target"],
                "Walker": ["torso_position", "torso_zaxis", "torso_subtreelinvel"],
                "Quadruped": ["torso_subtreecom", "torso_subtreelinvel", "torso_angmom"]}


def make_task(model, name: str) -> PsTaskC:
    t = PsTaskC()
    t.task_id = TASK_IDS[name]
    risk = model.numeric("task_risk")
    t.risk = float(risk[0]) if risk is not None else 0.0
    if model.nsensor == 0 or model.sensor_type[0] != P.SENS_USER:
        raise ValueError("user sensors specifying residuals must be specified first and sequentially")
    nterm = 1
    while nterm < model.nsensor and model.sensor_type[nterm] == P.SENS_USER:
        nterm += 1
    if nterm > MAX_COST_TERMS:
        raise ValueError("number of cost terms exceeds maximum")
    t.num_term = nterm
    nres, pshift = 0, 0
    for i in range(nterm):
        dim = int(model.sensor_dim[i]); nres += dim
        t.dim_norm_residual[i] = dim
        s = model.sensor_user[i]
        norm = int(s[0]); npar = NORM_PARAM_DIM[norm]
        if 4 + npar > model.nuser_sensor or any(s[4 + j] <= 0 for j in range(npar)):
            raise ValueError(f"cost construction: missing parameter value, sensor {i}")
        if norm == -1 and dim != 1:
            raise ValueError(f"cost construction: null norm needs dim 1, sensor {i}")
        t.norm[i] = norm; t.weight[i] = float(s[1]); t.num_norm_parameter[i] = npar
        for j in range(npar):
            t.norm_parameter[pshift + j] = float(s[4 + j])
        pshift += npar
    t.num_residual = nres
    ntrace = 0
    for i in range(model.nsensor):
        nm = model.name_of("sensor", i)
        if nm.startswith("trace"):
            if ntrace < 4:
                t.trace_sensor[ntrace] = model.name2id("sensor", f"trace{ntrace}")
            ntrace += 1
    t.num_trace = ntrace
    # residual_* numerics -> Task::parameters (SetFeatureParameters, task.cc:38-64)
    k = 0
    for i in range(model.nnumeric):
        nm = model.name_of("numeric", i)
        if nm.startswith("residual_select_"):
            # list selections are stored as ints reinterpreted as doubles (utilities.cc:207-211)
            v = np.array([int(model.numeric_data[model.numeric_adr[i]])], dtype=np.int64).view(np.float64)[0]
            t.parameters[k] = float(v); k += 1
        elif nm.startswith("residual_"):
            t.parameters[k] = float(model.numeric_data[model.numeric_adr[i]]); k += 1
    t.num_parameters = k
    for j in range(16):
        t.sensor_id[j] = -1
    for j, nm in enumerate(TASK_SENSORS.get(name, [])):
        sid = model.name2id("sensor", nm)
        if sid < 0:
            raise ValueError(f"task {name}: sensor '{nm}' not found in the model")
        t.sensor_id[j] = sid
    if name == "Quadruped":
        quadruped_reset(model, t)
        quadruped_snapshot(model, t, set_gait_parameters=False)   # state after the first Transition()
    return t


# --------------------------------------------------------------------------------------------
# Quadruped Flat: the ResidualFn member state the reference keeps between Transition() calls
# (mjpc/tasks/quadruped/quadruped.h:176-236) crosses the boundary as PsTask.aux / PsTask.iaux.
# Layout: mujoco_mpc_b200/csrc/ps_quadruped.cuh.
QUAD_MODES = {"Quadruped": 0, "Biped": 1, "Walk": 2, "Scramble": 3, "Flip": 4}
QUAD_GAITS = {"Stand": 0, "Walk": 1, "Trot": 2, "Canter": 3, "Gallop": 4}
# duty ratio, cadence, amplitude, balance, upright, height (quadruped.h:88-97)
QUAD_GAIT_PARAM = [[1, 1, 0, 0, 1, 1], [0.75, 1, 0.03, 0, 1, 1], [0.45, 2, 0.03, 0.2, 1, 1],
                   [0.4, 4, 0.05, 0.03, 0.5, 0.2], [0.3, 3.5, 0.10, 0.03, 0.2, 0.1]]
A_MODE, A_MODE_START, A_POSITION, A_HEADING, A_SPEED, A_ANGVEL, A_GAIT, A_PHASE_START, A_PHASE_START_TIME, \
    A_PHASE_VELOCITY, A_GROUND, A_ORIENTATION, A_FLIP, A_HOME, A_CROUCH = 0, 1, 2, 4, 6, 7, 8, 9, 10, 11, 12, 13, 20, 32, 51


def parameter_index(model, name: str) -> int:
    """ParameterIndex (mjpc/utilities.cc:178-196): position of residual_<name> among residual_* numerics."""
    k = 0
    for i in range(model.nnumeric):
        nm = model.name_of("numeric", i)
        if nm.startswith("residual_"):
            if nm == "residual_" + name:
                return k
            k += 1
    raise ValueError(f"parameter '{name}' not found")


def cost_term_by_name(model, name: str) -> int:
    """CostTermByName (mjpc/utilities.cc:241-250): index of the user sensor called `name`."""
    for i in range(model.nsensor):
        if model.sensor_type[i] == P.SENS_USER and model.name_of("sensor", i) == name:
            return i
    raise ValueError(f"cost term '{name}' not found")


def _as_double(i: int) -> float:
    return float(np.array([i], dtype=np.int64).view(np.float64)[0])


def quadruped_reset(model, t: PsTaskC) -> None:
    """QuadrupedFlat::ResetLocked (quadruped.cc:515-601): identifiers + flip trajectory constants,
    and the default-constructed ResidualFn state (mode Quadruped, gait Stand, phase clock at 0)."""
    for j in range(16):
        t.iaux[j] = -1
    for k in range(MAX_TASK_AUX):
        t.aux[k] = 0.0
    torso = model.name2id("body", "trunk"); head = model.name2id("site", "head"); goal = model.name2id("body", "goal")
    if torso < 0 or head < 0 or goal < 0:
        raise ValueError("Quadruped: body 'trunk', site 'head' and body 'goal' are required")
    mocap = int(model.body_mocapid[goal])
    if mocap < 0:
        raise ValueError("body 'goal' is not mocap")
    t.iaux[0], t.iaux[1], t.iaux[2] = torso, head, mocap
    for k, nm in enumerate(["FL", "HL", "FR", "HR"]):
        g = model.name2id("geom", nm)
        if g < 0:
            raise ValueError(f"geom '{nm}' not found")
        t.iaux[3 + k] = g
    t.iaux[7] = parameter_index(model, "select_Biped type")
    t.iaux[8] = parameter_index(model, "Heading")
    t.iaux[9] = parameter_index(model, "Amplitude")
    t.iaux[10] = parameter_index(model, "Duty ratio")
    t.iaux[11] = parameter_index(model, "select_Flip dir")
    # flip trajectory constants
    kHQ, kCrouch, kLeap, kMax = 0.25, 0.15, 0.5, 0.8
    gravity = float(np.linalg.norm(model.opt_gravity))
    jump_vel = np.sqrt(2 * gravity * (kMax - kLeap))
    flight_time = 2 * jump_vel / gravity
    jump_acc = jump_vel * jump_vel / (2 * (kLeap - kCrouch))
    crouch_time = np.sqrt(2 * (kHQ - kCrouch) / jump_acc)
    leap_time = jump_vel / jump_acc
    jump_time = crouch_time + leap_time
    crouch_vel = -jump_acc * crouch_time
    land_time = 2 * (kLeap - kHQ) / jump_vel
    land_acc = jump_vel / land_time
    flight_rot_vel = 1.25 * np.pi / flight_time
    jump_rot_vel = np.pi / leap_time - flight_rot_vel
    jump_rot_acc = (flight_rot_vel - jump_rot_vel) / leap_time
    land_rot_acc = 2 * (flight_rot_vel * land_time - np.pi / 4) / (land_time * land_time)
    for k, v in enumerate([crouch_time, jump_time, flight_time, land_time, crouch_vel, jump_acc, jump_vel, land_acc,
                           jump_rot_vel, jump_rot_acc, flight_rot_vel, land_rot_acc]):
        t.aux[A_FLIP + k] = float(v)
    for key, base in (("home", A_HOME), ("crouch", A_CROUCH)):
        kid = model.name2id("key", key)
        if kid < 0:
            raise ValueError(f"keyframe '{key}' not found")
        for k in range(model.nq):
            t.aux[base + k] = float(np.asarray(model.key_qpos).reshape(-1, model.nq)[kid, k])


def quadruped_snapshot(model, t: PsTaskC, *, mode="Quadruped", gait="Stand", time=0.0, mode_start_time=None,
                       position=(0.0, 0.0), heading=(0.0, 0.0), speed=0.0, angvel=0.0, phase_start=0.0,
                       phase_start_time=None, ground=0.0, orientation=(1.0, 0.0, 0.0, 0.0), set_gait_parameters=True) -> None:
    """State QuadrupedFlat::TransitionLocked (quadruped.cc:224-388) would have left in the ResidualFn when
    the planner snapshots it: the mode, the gait with its duty / cadence / amplitude parameters and
    balance / upright / height weights (:283-293), the phase clock (:247-252) and the Walk / Flip anchors."""
    g = QUAD_GAITS[gait] if isinstance(gait, str) else int(gait)
    md = QUAD_MODES[mode] if isinstance(mode, str) else int(mode)
    if md == QUAD_MODES["Biped"]:
        g = QUAD_GAITS["Trot"]
    if set_gait_parameters:
        gp = QUAD_GAIT_PARAM[g]
        t.parameters[parameter_index(model, "select_Gait")] = _as_double(g)
        t.parameters[parameter_index(model, "Duty ratio")] = gp[0]
        t.parameters[parameter_index(model, "Cadence")] = gp[1]
        t.parameters[parameter_index(model, "Amplitude")] = gp[2]
        t.weight[cost_term_by_name(model, "Balance")] = gp[3]
        t.weight[cost_term_by_name(model, "Upright")] = gp[4]
        t.weight[cost_term_by_name(model, "Height")] = gp[5]
    cadence = t.parameters[parameter_index(model, "Cadence")]
    t.aux[A_MODE] = float(md)
    t.aux[A_MODE_START] = float(time if mode_start_time is None else mode_start_time)
    t.aux[A_POSITION], t.aux[A_POSITION + 1] = float(position[0]), float(position[1])
    t.aux[A_HEADING], t.aux[A_HEADING + 1] = float(heading[0]), float(heading[1])
    t.aux[A_SPEED], t.aux[A_ANGVEL], t.aux[A_GAIT] = float(speed), float(angvel), float(g)
    t.aux[A_PHASE_START] = float(phase_start)
    t.aux[A_PHASE_START_TIME] = float(time if phase_start_time is None else phase_start_time)
    t.aux[A_PHASE_VELOCITY] = 2 * np.pi * cadence
    t.aux[A_GROUND] = float(ground)
    for k in range(4):
        t.aux[A_ORIENTATION + k] = float(orientation[k])
