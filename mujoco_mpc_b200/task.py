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
                "Walker": ["torso_position", "torso_zaxis", "torso_subtreelinvel"]}


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
    return t
