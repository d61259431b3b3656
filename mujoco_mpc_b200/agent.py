"""In-process mirror of the reference's Python client `mujoco_mpc.agent.Agent` (python/mujoco_mpc/agent.py:68-400)
and of the service calls behind it (mjpc/grpc/grpc_agent_util.cc): same method names and argument meaning, without
the gRPC transport -- the planner is the B200 engine (CudaSamplingPlanner / CudaCrossEntropyPlanner /
CudaRobustPlanner) and `step()` runs the plant's mj_step on the device as well.

    with Agent("Cartpole") as agent:
        agent.set_state(qpos=..., qvel=..., time=0.0)
        agent.planner_step()
        u = agent.get_action()

Plumbing around the C ABI, like testspeed.py; nothing here is on the planning hot path.
"""
import math

import numpy as np

from . import registry, task as _task
from .engine import Engine
from .planner import CudaCrossEntropyPlanner, CudaRobustPlanner, CudaSamplingPlanner
from .testspeed import plant_model, transition_mocap

_PLANNERS = {"Sampling": CudaSamplingPlanner, "Cross Entropy": CudaCrossEntropyPlanner, "Robust Sampling": CudaRobustPlanner}


def norm_value(kind, x, params=()):
    """mjpc::Norm (mjpc/norm.cc:50-210), value only: the per-term cost of a residual block."""
    x = np.asarray(x, dtype=np.float64)
    p = params[0] if len(params) > 0 else 0.0
    q = params[1] if len(params) > 1 else 0.0
    if kind == -1:                                    # kNull
        return float(x[0])
    if kind == 0:                                     # kQuadratic
        return 0.5 * float(np.dot(x, x))
    if kind == 1:                                     # kL22
        c = float(np.dot(x, x))
        return (c ** (q / 2) + p ** q) ** (1 / q) - p
    if kind == 2:                                     # kL2
        return math.sqrt(float(np.dot(x, x)) + p * p) - p
    if kind == 3:                                     # kCosh
        return float(np.sum(p * p * (np.cosh(x / p) - 1.0)))
    if kind == 5:                                     # kPowerLoss
        return float(np.sum(np.abs(x) ** p))
    if kind == 6:                                     # kSmoothAbsLoss
        return float(np.sum(np.sqrt(x * x + p * p) - p))
    if kind == 7:                                     # kSmoothAbs2Loss
        return float(np.sum((np.abs(x) ** q + p ** q) ** (1 / q) - p))
    if kind == 8:                                     # kRectifyLoss
        return float(np.sum(p * np.log1p(np.exp(x / p)))) if p > 0 else float(np.sum(np.maximum(x, 0)))
    raise ValueError(f"norm type {kind}")


class Agent:
    def __init__(self, task_id, device=0, planner="Sampling"):
        if planner not in _PLANNERS:
            raise ValueError(f"planner {planner!r}: one of {sorted(_PLANNERS)}")
        self.task_id = task_id
        self.model, self.task, self.settings = registry.load(task_id)             # the agent's copy: agent_timestep / integrator
        self.plant_model = plant_model(task_id)                                   # what Step() integrates: the XML's own options
        self.plant_task = _task.make_task(self.plant_model, registry._TASK_ID_NAME.get(task_id, task_id))
        self.planner = _PLANNERS[planner](self.model, self.task, device=device)
        self.plant = Engine(self.plant_model, self.plant_task, device=device)
        m = self.model
        self._term_names = [m.name_of("sensor", i) for i in range(self.task.num_term)]
        self._param_names = [m.name_of("numeric", i) for i in range(m.nnumeric) if m.name_of("numeric", i).startswith("residual_")]
        tr = m.text("task_transition")                      # Task::mode names (grpc_agent_util.cc GetAllModes)
        self._modes = tr.split("|") if tr else ["default_mode"]
        self._mode = 0
        self._default_weights = [self.task.weight[i] for i in range(self.task.num_term)]
        self._default_parameters = [self.task.parameters[i] for i in range(self.task.num_parameters)]
        self.reset()

    # ---- context manager / lifetime (agent.py:120-133)
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        for x in (getattr(self, "planner", None), getattr(self, "plant", None)):
            if x is not None:
                x.close()
        self.planner = self.plant = None

    # ---- state (SetState / GetState, grpc_agent_util.cc)
    def reset(self):
        """Reset: data to the "home" keyframe (or qpos0), planner to a zero plan, task weights to the model's."""
        pm = self.plant_model
        st = registry.home_state(pm)
        self.time = 0.0
        self.qpos, self.qvel, self.act = st[:pm.nq].copy(), st[pm.nq:pm.nq + pm.nv].copy(), st[pm.nq + pm.nv:].copy()
        mo = transition_mocap(None, pm, 0.0)
        self.mocap = mo.reshape(-1, 7).copy() if mo is not None else np.zeros((0, 7))
        self.userdata = np.zeros(pm.nuserdata)
        self.ctrl = np.zeros(pm.nu)
        self._warmstart = np.zeros(pm.nv)              # mjData::qacc_warmstart of the agent-side plant (mj_resetData: zeros)
        for i, w in enumerate(self._default_weights):
            self.task.weight[i] = w; self.plant_task.weight[i] = w
        for i, v in enumerate(self._default_parameters):           # Task::Reset re-reads the residual_* numerics
            self.task.parameters[i] = v; self.plant_task.parameters[i] = v
        self.planner.Reset(self.settings["steps"], np.zeros(self.model.nu))
        self._push_task()

    def set_state(self, time=None, qpos=None, qvel=None, act=None, mocap_pos=None, mocap_quat=None, userdata=None):
        def put(dst, src, what):
            src = np.asarray(src, dtype=np.float64).reshape(-1)
            if src.size != dst.size:
                raise ValueError(f"{what}: expected {dst.size} values, got {src.size}")       # INVALID_ARGUMENT in the service
            dst[...] = src.reshape(dst.shape)
        if time is not None:
            self.time = float(time)
        if qpos is not None and np.size(qpos):
            put(self.qpos, qpos, "qpos")
        if qvel is not None and np.size(qvel):
            put(self.qvel, qvel, "qvel")
        if act is not None and np.size(act):
            put(self.act, act, "act")
        if mocap_pos is not None and np.size(mocap_pos):
            put(self.mocap[:, :3], mocap_pos, "mocap_pos")
        if mocap_quat is not None and np.size(mocap_quat):
            put(self.mocap[:, 3:], mocap_quat, "mocap_quat")
        if userdata is not None and np.size(userdata):
            put(self.userdata, userdata, "userdata")

    def get_state(self):
        return dict(time=self.time, qpos=self.qpos.copy(), qvel=self.qvel.copy(), act=self.act.copy(),
                    mocap_pos=self.mocap[:, :3].reshape(-1).copy(), mocap_quat=self.mocap[:, 3:].reshape(-1).copy(),
                    userdata=self.userdata.copy())

    def set_mocap(self, mocap_map):
        """{mocap body name: (pos or None, quat or None)}"""
        pm = self.plant_model
        for name, (pos, quat) in mocap_map.items():
            b = pm.name2id("body", name)
            if b < 0 or pm.body_mocapid[b] < 0:
                raise ValueError(f"Body '{name}' is not a mocap body")
            i = int(pm.body_mocapid[b])
            if pos is not None:
                self.mocap[i, :3] = pos
            if quat is not None:
                q = np.asarray(quat, dtype=np.float64)
                self.mocap[i, 3:] = q / np.linalg.norm(q)          # the service normalises (agent_test.py: test_set_mocap)

    def _state_vector(self):
        return np.concatenate([self.qpos, self.qvel, self.act])

    def _mocap_flat(self):
        return self.mocap.reshape(-1) if self.mocap.size else None

    # ---- planning (PlannerStep -> Agent::PlanIteration, agent.cc:270-315)
    def _push_task(self):
        self.planner.SetTask(self.task)
        self.plant.set_task(self.plant_task)

    def planner_step(self):
        self.planner.SetState(self._state_vector(), self.time, mocap=self._mocap_flat())
        self.planner.OptimizePolicy(self.settings["steps"])

    def get_action(self, time=None, averaging_duration=0, nominal_action=False):
        """GetAction / AverageAction (grpc_agent_util.cc:166-230).  The sampling policies are open loop (the state
        argument of ActionFromPolicy is unused), so the averaged action needs no rollout: the policy is sampled on
        the grid the service walks -- `< end` for the nominal action, `<= end` otherwise.  The step of that grid is
        model->opt.timestep of the SERVICE's model copy, taken before the agent overrides timestep / integrator
        (grpc/agent_service.cc:118-120), i.e. the XML plant time step (Cartpole 0.001, not the agent's 0.01)."""
        t = self.time if time is None else float(time)
        if averaging_duration > 0:
            end, dt = t + averaging_duration, self.plant_model.opt_timestep
            total, n = np.zeros(self.model.nu), 0
            while (t < end) if nominal_action else (t <= end):
                total += self.planner.ActionFromPolicy(t); t += dt; n += 1
            return total / max(n, 1)
        return self.planner.ActionFromPolicy(t)

    def best_trajectory(self):
        bt = self.planner.BestTrajectory()
        if bt is None:
            return None
        return dict(states=bt["states"], actions=bt["actions"][:-1], times=bt["times"])      # agent.py:368-383

    # ---- physics on the agent side (Step: Transition, action, mj_step)
    def _forward_record(self, steps):
        """`steps` plant steps from the current state under the planner's policy; the writeback record."""
        times, knots = self.planner.policy()
        if len(times) == 0:
            times, knots = np.zeros(1), np.zeros((1, self.model.nu))
        interp = self.settings["interpolation"]
        self.plant.set_state(self._state_vector(), self.time, mocap=self._mocap_flat())
        self.plant.set_nominal(times, knots, interp)
        self.plant.set_warmstart(self._warmstart)
        self.plant.rollouts(1, steps + 1, 0.0, writeback=True)
        return self.plant.trajectory(0)

    def step(self):
        mo = transition_mocap(self.task_id, self.plant_model, self.time)
        if mo is not None and self.task_id == "Particle":
            self.mocap = mo.reshape(-1, 7).copy()
        tr = self._forward_record(1)
        self._warmstart = self.plant.warmstart(0)      # mj_step leaves its solution in qacc_warmstart for the next step
        pm = self.plant_model
        self.ctrl = tr["actions"][0].copy()
        st = tr["states"][1]
        self.qpos, self.qvel, self.act = st[:pm.nq].copy(), st[pm.nq:pm.nq + pm.nv].copy(), st[pm.nq + pm.nv:].copy()
        self.time = float(tr["times"][1])

    # ---- costs (GetCostValuesAndWeights / GetResiduals)
    def _residual_now(self):
        return self._forward_record(1)["residual"][0]

    def get_residuals(self):
        r, out, k = self._residual_now(), {}, 0
        for i, name in enumerate(self._term_names):
            d = self.task.dim_norm_residual[i]
            out[name] = r[k:k + d].copy(); k += d
        return out

    def get_cost_term_values(self):
        r, out, k, pshift = self._residual_now(), {}, 0, 0
        t = self.plant_task
        for i, name in enumerate(self._term_names):
            d, npar = t.dim_norm_residual[i], t.num_norm_parameter[i]
            out[name] = norm_value(t.norm[i], r[k:k + d], [t.norm_parameter[pshift + j] for j in range(npar)])
            k += d; pshift += npar
        return out

    def get_total_cost(self):
        w = self.get_cost_weights()
        return sum(w[name] * v for name, v in self.get_cost_term_values().items())

    def get_cost_weights(self):
        return {name: self.task.weight[i] for i, name in enumerate(self._term_names)}

    def set_cost_weights(self, weights, reset_to_defaults=False):
        if reset_to_defaults:
            for i, w in enumerate(self._default_weights):
                self.task.weight[i] = w
        for name, w in weights.items():
            if name not in self._term_names:
                raise ValueError(f"Weight '{name}' not found in task '{self.task_id}'")
            self.task.weight[self._term_names.index(name)] = float(w)
        for i in range(self.task.num_term):
            self.plant_task.weight[i] = self.task.weight[i]
        self._push_task()

    # ---- task parameters (residual_<name> numerics, task.cc:38-64)
    def get_task_parameters(self):
        return {n[len("residual_"):]: self.task.parameters[i] for i, n in enumerate(self._param_names)
                if not n.startswith("residual_select_")}

    def set_task_parameter(self, name, value):
        self.set_task_parameters({name: value})

    def set_task_parameters(self, parameters):
        for name, value in parameters.items():
            key = "residual_" + name
            if key not in self._param_names:
                raise ValueError(f"Parameter '{name}' not found in task '{self.task_id}'")
            i = self._param_names.index(key)
            self.task.parameters[i] = float(value); self.plant_task.parameters[i] = float(value)
        self._push_task()

    # ---- modes (Task::mode; custom text "task_transition")
    def get_all_modes(self):
        return list(self._modes)

    def get_mode(self):
        return self._modes[self._mode]

    def set_mode(self, mode):
        if mode not in self._modes:
            raise ValueError(f"Mode '{mode}' not found in task '{self.task_id}'")       # INVALID_ARGUMENT in the service
        self._mode = self._modes.index(mode)
        if self.task_id.startswith("Quadruped"):
            # what the task's next Transition() leaves in the residual snapshot: the mode and its start time
            for mdl, tsk in ((self.model, self.task), (self.plant_model, self.plant_task)):
                _task.quadruped_snapshot(mdl, tsk, mode=mode, time=self.time, set_gait_parameters=False)
            self._push_task()
