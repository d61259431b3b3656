"""Shared helpers for the parity tests: seeded synthetic inputs and the
oracle-side planning iteration with injected noise."""
import numpy as np
from mujoco_mpc_b200 import registry
from oracle import pyoracle


from mujoco_mpc_b200.registry import synthetic_state  # noqa: F401


def oracle_iteration(model, task, settings, state, time, N, H, P, interp, sigma, z, nominal=None, mocap=None, nthread=1):
    """One OptimizePolicy on the oracle.  Returns (oracle, resampled nominal times, knots, winner)."""
    o = pyoracle.Oracle(model, task)
    o.configure(P, interp, sigma)
    o.set_state(state, time, mocap=mocap)
    if nominal is not None:
        o.set_policy(nominal[0], nominal[1], interp)
    w = o.optimize(N, H, z, nthread=nthread)
    nt, nk = o.policy(1)      # previous_policy = the resampled nominal of this iteration
    return o, nt, nk, w


def model_mocap(model):
    """mocap pose [nmocap][7] of the model as loaded (what mjData holds after mj_resetData)"""
    out = np.zeros((max(model.nmocap, 0), 7))
    for b in range(model.nbody):
        i = int(model.body_mocapid[b])
        if i >= 0:
            out[i, :3] = np.asarray(model.body_pos).reshape(-1, 3)[b]
            out[i, 3:] = np.asarray(model.body_quat).reshape(-1, 4)[b]
    return out.reshape(-1) if model.nmocap else None


def run_isolated(module, func, *args, timeout=900):
    """Run `module.func(*args)` in a child interpreter.  Used by GPU tests whose CUDA path has not run on hardware
    yet: a device fault or an abort() in there must not take the verified tests of the same session down with it."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = f"import {module} as m; m.{func}(*{args!r}); print('ISOLATED-OK')"
    r = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0 and "ISOLATED-OK" in r.stdout, f"exit {r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"


def oracle_closed_loop(task_name, spp, total_time, N, interp, noise):
    """mjpc/testspeed.cc:95-113 restated on the oracle: Transition; agent.state.Set (the state BEFORE the step);
    action; mj_step; cost; after every spp-th step PlanIteration from agent.state (agent.cc:270-315)."""
    import math
    from mujoco_mpc_b200 import task as T
    from mujoco_mpc_b200.testspeed import plant_model, transition_mocap
    model, task, s = registry.load(task_name)
    pm = plant_model(task_name)
    ptask = T.make_task(pm, registry._TASK_ID_NAME.get(task_name, task_name))
    P = min(s["num_spline_points"], 36)
    planner = pyoracle.Oracle(model, task); planner.configure(P, interp, s["exploration"])
    planner.set_policy(np.zeros(1), np.zeros((1, model.nu)), interp)       # agent.Reset(data->ctrl)
    plant = pyoracle.Oracle(pm, ptask)
    st = registry.home_state(pm)
    plant.reset_data(); plant.set("qpos", st[:pm.nq]); plant.set("qvel", st[pm.nq:pm.nq + pm.nv]); plant.set("time", [0.0])
    total_steps = int(math.ceil(total_time / pm.opt_timestep)); total = 0.0; plans = 0

    def plant_state():
        return np.concatenate([plant.get("qpos")[:pm.nq], plant.get("qvel")[:pm.nv], plant.get("act")[:pm.na]])
    for i in range(total_steps):
        t = plant.get("time")[0]
        mo = transition_mocap(task_name, pm, t)                 # Task::Transition
        if mo is not None:
            plant.set("mocap_pos", mo.reshape(-1, 7)[:, :3].reshape(-1)); plant.set("mocap_quat", mo.reshape(-1, 7)[:, 3:].reshape(-1))
        agent_state, agent_time = plant_state(), t              # agent.state.Set(model, data)
        plant.set("ctrl", planner.action(t))
        plant.step()
        total += plant.cost(plant.get("sensordata")[:ptask.num_residual])
        if i % spp == 0:
            planner.set_state(agent_state, agent_time, mocap=mo)
            planner.optimize(N, s["steps"], noise(plans, N, P, model.nu)); plans += 1
    return dict(total_cost=total, state=plant_state(), time=plant.get("time")[0], plans=plans, policy=planner.policy(0))
