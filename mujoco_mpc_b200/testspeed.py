"""Headless closed loop of the reference's `testspeed` harness (mjpc/testspeed.cc:43-132,
`SynchronousPlanningCost`) with BOTH sides of the loop on the B200: the planner is
CudaSamplingPlanner, and the plant's `mj_step` is the same device forward step, run as a one-sample
rollout with writeback between planning iterations (BASELINE.json
configs[0] is `testspeed --task=Cartpole` semantics).

Loop, per plant step i (testspeed.cc:95-113): Transition; agent.state.Set(data) -- the state BEFORE the step;
action = ActionFromPolicy(state, time); mj_step; cost = CostValue(sensordata); after every
`steps_per_planning_iteration`-th step PlanIteration (agent.cc:270-315: planner.SetState(agent.state) -- still
the pre-step state --, steps = clamp(horizon / agent_timestep + 1, 1, 512), OptimizePolicy).  Covered: the tasks
whose Transition() is the base no-op (cartpole, acrobot, swimmer, walker) and Particle (Transition moves the goal
mocap along a Lissajous curve, particle.cc:58-65).  The plant runs as one launch per planning period; the
constraint solver's warm start is handed from one launch to the next (ps_get_warmstart / ps_set_warmstart), as the
reference's one mjData carries qacc_warmstart from step to step.  Quadruped Flat runs with the mode / gait of its
reset state; its Transition (gait switching, quadruped.cc) is not in this loop.

    python -m mujoco_mpc_b200.testspeed --task Cartpole --steps_per_planning_iteration 4 --total_time 10
"""
import argparse
import math
import os
import time as _time

import numpy as np

from . import mjcf, registry, task as _task
from .engine import Engine
from .planner import CudaSamplingPlanner


def plant_model(name):
    """The model as the plant sees it: the XML's own opt.timestep / integrator (the agent overrides only
    touch the Agent's private copy, agent.cc:74,275-276)."""
    return mjcf.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tasks", registry.TASKS[name]))


def _model_mocap(model):
    out = np.zeros((max(model.nmocap, 1), 7))
    for b in range(model.nbody):
        i = int(model.body_mocapid[b])
        if i >= 0:
            out[i, :3] = np.asarray(model.body_pos).reshape(-1, 3)[b]
            out[i, 3:] = np.asarray(model.body_quat).reshape(-1, 4)[b]
    return out


def transition_mocap(task_name, model, time):
    """Task::Transition() as far as it reaches the planner: the mocap poses [nmocap][7] at `time` (None: no mocap)."""
    if model.nmocap == 0:
        return None
    mo = _model_mocap(model)
    if task_name == "Particle":                      # particle.cc:58-65: x, y of the goal follow a Lissajous curve
        mo[0, 0] = 0.25 * math.sin(time); mo[0, 1] = 0.25 * math.cos(time / math.pi)
    return mo.reshape(-1)


def synchronous_planning_cost(task_name, steps_per_planning_iteration=4, total_time=10.0, num_trajectory=None,
                              interpolation=None, seed=1, noise=None, device=0, verbose=False):
    """Returns dict(total_cost, average_cost, steps, plans, wall_s, plan_s, state, time).  `noise(k, N, P, nu)` may
    return the injected N(0,1) tensor of planning iteration k (parity runs); default is the device generator."""
    model, task, s = registry.load(task_name)
    pm = plant_model(task_name)
    ptask = _task.make_task(pm, registry._TASK_ID_NAME.get(task_name, task_name))
    N = num_trajectory or s["num_trajectory"]
    interp = s["interpolation"] if interpolation is None else interpolation
    P = min(s["num_spline_points"], 36)
    pl = CudaSamplingPlanner(model, task, device=device)
    pl.configure(num_trajectory=N, interpolation=interp, sigma0=s["exploration"], sigma1=s["exploration2"])
    pl.SetSeed(seed)
    plant = Engine(pm, ptask, device=device)
    state = registry.home_state(pm)                 # mj_resetDataKeyframe(model, data, "home"), testspeed.cc:72-76
    pl.Reset(s["steps"], np.zeros(model.nu))          # agent.Reset(data->ctrl)
    t = 0.0
    total_steps = int(math.ceil(total_time / pm.opt_timestep))
    total_cost, plans, plan_s = 0.0, 0, 0.0
    t0 = _time.perf_counter()
    i = 0
    while i < total_steps:
        # plant steps until (and including) the next step after which the planner runs
        spp = steps_per_planning_iteration
        k = min(-(-i // spp) * spp - i + 1, total_steps - i)      # the planner runs after steps 0, spp, 2 spp, ...
        times, knots = pl.policy()
        if len(times) == 0:
            times, knots = np.zeros(1), np.zeros((1, model.nu))
        plant.set_state(state, t, mocap=transition_mocap(task_name, pm, t))
        plant.set_nominal(times, knots, interp)
        plant.rollouts(1, k + 1, 0.0, writeback=True)
        tr = plant.trajectory(0)
        plant.set_warmstart(plant.warmstart(0))        # one mjData stepped by mj_step carries qacc_warmstart over
        total_cost += float(np.sum(tr["costs"][:k]))
        # agent.state was captured BEFORE the last step (testspeed.cc:97 precedes mj_step): that is what plans
        plan_state = tr["states"][k - 1].copy(); plan_time = float(tr["times"][k - 1])
        state = tr["states"][k].copy(); t = float(tr["times"][k])
        last = i + k - 1
        i += k
        if last % steps_per_planning_iteration == 0:
            p0 = _time.perf_counter()
            if noise is not None:
                pl.SetNoise(noise(plans, N, P, model.nu))
            pl.SetState(plan_state, plan_time, mocap=transition_mocap(task_name, model, plan_time))
            pl.OptimizePolicy(s["steps"])
            plan_s += _time.perf_counter() - p0
            plans += 1
        if verbose and math.floor(t) > math.floor(t - k * pm.opt_timestep):
            print(f"sim time: {math.floor(t)}, cost: {tr['costs'][k - 1]:.6g}")
    wall = _time.perf_counter() - t0
    pl.close(); plant.close()
    return dict(total_cost=total_cost, average_cost=total_cost / total_steps, steps=total_steps, plans=plans, wall_s=wall,
                plan_s=plan_s, state=state, time=t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Cartpole")
    ap.add_argument("--steps_per_planning_iteration", type=int, default=4)
    ap.add_argument("--total_time", type=float, default=10.0)
    ap.add_argument("--num_trajectory", type=int, default=None)
    ap.add_argument("--interpolation", type=int, default=None, help="0 zero, 1 linear, 2 cubic (default: the task's)")
    a = ap.parse_args()
    r = synchronous_planning_cost(a.task, a.steps_per_planning_iteration, a.total_time, a.num_trajectory, a.interpolation, verbose=True)
    print(f"Total wall time ({r['plans']} planning steps): {r['wall_s']:.3f} s ({a.total_time / r['wall_s']:.2f}x realtime); "
          f"planning {r['plan_s'] / max(r['plans'], 1) * 1e3:.3f} ms / iteration")
    print(f"Average cost per step (lower is better): {r['average_cost']:.6g}")


if __name__ == "__main__":
    main()
