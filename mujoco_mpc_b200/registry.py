"""Task registry: model XML + reference task name + the agent / planner settings the reference reads
from the model's custom numerics (mjpc/agent.cc:90-107, planners/sampling/planner.cc:53-67,
planners/sampling/policy.cc:36-37)."""
import os
from . import mjcf, task as _task
from . import psmodel as P

_HERE = os.path.dirname(os.path.abspath(__file__))
TASKS = {  # reference Task::Name() -> model file
    "Cartpole": "cartpole.xml", "Acrobot": "acrobot.xml", "Particle": "particle.xml",
    "Swimmer": "swimmer.xml", "Walker": "walker.xml", "Quadruped Flat": "quadruped_flat.xml",
}
_TASK_ID_NAME = {"Quadruped Flat": "Quadruped"}
K_MAX_TRAJECTORY_HORIZON = 512   # mjpc/trajectory.h:27


def _num(model, name, default):
    v = model.numeric(name)
    return default if v is None else float(v[0])


def load(name):
    """Returns (model, task, settings).  The model's opt.timestep / opt.integrator already carry
    the agent overrides Agent::PlanIteration writes every iteration (agent.cc:275-276)."""
    path = os.path.join(_HERE, "tasks", TASKS[name])
    if not os.path.exists(path):
        raise FileNotFoundError(f"task model {path} not available")
    m = mjcf.load(path)
    s = dict(
        horizon=_num(m, "agent_horizon", 0.5), timestep=_num(m, "agent_timestep", 1.0e-2),
        integrator=int(_num(m, "agent_integrator", m.opt_integrator)),
        num_spline_points=int(_num(m, "sampling_spline_points", K_MAX_TRAJECTORY_HORIZON)),
        interpolation=int(_num(m, "sampling_representation", _task.SPLINE_CUBIC)),
        num_trajectory=int(_num(m, "sampling_trajectories", 10)),
        exploration=_num(m, "sampling_exploration", 0.1), exploration2=0.0)
    se = m.numeric("sampling_exploration")
    if se is not None and len(se) > 1:
        s["exploration2"] = float(se[1])
    s["steps"] = int(max(min(s["horizon"] / s["timestep"] + 1, K_MAX_TRAJECTORY_HORIZON), 1))
    m.v["opt_timestep"] = s["timestep"]
    m.v["opt_integrator"] = s["integrator"]
    m._c = None
    t = _task.make_task(m, _TASK_ID_NAME.get(name, name))
    return m, t, s


def home_state(model):
    """qpos of keyframe 'home' (testspeed.cc:72-76) or qpos0, zero velocity / activation."""
    import numpy as np
    st = np.zeros(model.nq + model.nv + model.na)
    k = model.name2id("key", "home")
    st[:model.nq] = model.key_qpos[k] if k >= 0 else model.qpos0
    return st


def synthetic_state(model, seed=1234, pos=0.05, vel=0.1):
    import numpy as np
    rng = np.random.default_rng(seed)
    st = home_state(model)
    st[:model.nq] += rng.uniform(-pos, pos, model.nq)
    st[model.nq:model.nq + model.nv] += rng.uniform(-vel, vel, model.nv)
    for j in range(model.njnt):        # renormalise free-joint quaternions
        if model.jnt_type[j] == 0:
            a = model.jnt_qposadr[j] + 3
            st[a:a + 4] /= np.linalg.norm(st[a:a + 4])
    return st
