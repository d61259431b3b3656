"""mujoco_mpc_b200/agent.py -- the in-process mirror of the reference's Python `Agent` client -- against the cases of
python/mujoco_mpc/agent_test.py, re-expressed.  CPU: planner and plant run on the host stand-in of the engine
(tests/hostsim/hostengine.cc), which has no noise generator, so every planner step gets numpy noise injected."""
import ctypes
import json
import os
import numpy as np
import pytest
from mujoco_mpc_b200 import engine, planner
from tests.hostsim import driver


@pytest.fixture()
def hostlib(oracle_lib, monkeypatch):
    L = planner.prototype(ctypes.CDLL(driver.build_planner_host()))
    monkeypatch.setattr(planner, "_LIB", L)
    monkeypatch.setattr(engine, "_LIB", engine.prototype(L))
    return L


def make(task_id, **kw):
    from mujoco_mpc_b200.agent import Agent
    a = Agent(task_id, **kw)
    a._rng = np.random.default_rng(7)
    return a


def plan(agent, N=None):
    s = agent.settings
    N = N or s["num_trajectory"]
    P = min(s["num_spline_points"], 36)
    d = agent.planner.delegate if hasattr(agent.planner, "delegate") else agent.planner
    d.configure(num_trajectory=N)
    d.SetNoise(agent._rng.standard_normal((N, P, agent.model.nu)))
    agent.planner_step()


def test_norms_against_the_reference_golden_vectors():
    # the numpy norms of agent.py (cost term values) against mjpc/norm.cc compiled from the reference (tests/golden)
    from mujoco_mpc_b200.agent import norm_value
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "norm_spline_reference.json")))
    cases = [c for c in g["norm"]] if isinstance(g, dict) and "norm" in g else []
    assert cases, "golden file layout changed"
    for c in cases:
        v = norm_value(int(c["type"]), c["x"], c["params"])
        assert abs(v - c["value"]) <= 1e-12 * max(1.0, abs(c["value"])), c


def test_set_task_parameters(hostlib):
    with make("Cartpole") as agent:
        agent.set_task_parameters({"Goal": 13})
        assert agent.get_task_parameters()["Goal"] == 13
        agent.set_task_parameter("Goal", 14)
        assert agent.get_task_parameters()["Goal"] == 14
        with pytest.raises(ValueError):
            agent.set_task_parameter("No such parameter", 1.0)
        agent.reset()
        assert agent.get_task_parameters()["Goal"] == agent._default_parameters[agent._param_names.index("residual_Goal")]


def test_step_env_with_planner_and_best_trajectory(hostlib):
    with make("Particle") as agent:
        actions, observations = [], []
        for _ in range(5):
            plan(agent, N=8)
            actions.append(agent.get_action())
            agent.step()
            observations.append(np.concatenate([agent.qpos, agent.qvel]))
        assert not (np.array(observations) == 0).all() and not (np.array(actions) == 0).all()
        bt = agent.best_trajectory()
        assert bt["states"].shape == (51, 4) and bt["actions"].shape == (50, 2) and bt["times"].shape == (51,)


def test_env_initialized_to_home_keyframe(hostlib):
    with make("Cartpole") as agent:
        st = agent.get_state()
        pm = agent.plant_model
        k = pm.name2id("key", "home")
        expect = pm.key_qpos[k] if k >= 0 else pm.qpos0
        assert np.allclose(st["qpos"], expect) and st["time"] == 0.0


def test_stepping_on_agent_side(hostlib):
    with make("Cartpole") as agent:
        agent.set_task_parameter("Goal", -1.0)
        obs = []
        for _ in range(4):
            plan(agent, N=8); agent.step()
            st = agent.get_state(); obs.append(np.concatenate([st["qpos"], st["qvel"]]))
        assert agent.get_state()["time"] == pytest.approx(4 * agent.plant_model.opt_timestep)
        assert not (np.array(obs) == 0).all()


def test_set_cost_weights_zero_action(hostlib):
    with make("Cartpole") as agent:
        agent.set_task_parameter("Goal", -1.0)
        plan(agent, N=64)
        assert not np.allclose(agent.get_action(), 0)
        # every cost but the control cost switched off: the best plan is to do nothing
        agent.reset()
        agent.set_task_parameter("Goal", -1.0)
        agent.set_cost_weights({"Vertical": 0, "Velocity": 0, "Centered": 0, "Control": 1})
        plan(agent, N=64)
        assert np.allclose(agent.get_action(), 0)
        with pytest.raises(ValueError):
            agent.set_cost_weights({"No such term": 1})


def test_get_cost_weights_and_total_cost(hostlib):
    with make("Cartpole") as agent:
        agent.set_task_parameter("Goal", -1.0)
        plan(agent, N=8)
        assert agent.get_total_cost() != 0
        terms, res = agent.get_cost_term_values(), agent.get_residuals()
        assert set(terms) == set(res) == {"Vertical", "Centered", "Velocity", "Control"}
        assert terms["Centered"] > 1e-3 and np.allclose(res["Centered"], [agent.qpos[0] + 1.0])       # cart at 0, goal at -1
        default = agent.get_cost_weights()
        agent.set_cost_weights({"Vertical": 1, "Velocity": 0, "Centered": 1, "Control": 0})
        assert agent.get_cost_weights() == {"Vertical": 1, "Velocity": 0, "Centered": 1, "Control": 0}
        agent.set_cost_weights({"Velocity": 0.5}, reset_to_defaults=True)
        assert agent.get_cost_weights() == {**default, "Velocity": 0.5}
        # total cost = sum of weight x term value, and equals what the device computes for the same state
        total = agent.get_total_cost()
        tr = agent._forward_record(1)
        assert total == pytest.approx(tr["costs"][0], rel=1e-12)


def test_action_averaging(hostlib):
    with make("Cartpole") as agent:
        agent.set_task_parameter("Goal", 0.5)
        plan(agent, N=8)
        # the averaging grid is the PLANT time step (0.001 for Cartpole), not the agent's 0.01 (agent_service.cc:118-120)
        dt = agent.plant_model.opt_timestep
        assert dt != agent.settings["timestep"]
        a0 = agent.get_action(time=0.0)
        avg = agent.get_action(time=0.0, averaging_duration=3 * dt)
        grid = [agent.get_action(time=k * dt) for k in range(5)]
        # `<= end` walks t = 0, dt, 2 dt, 3 dt (and one more when rounding leaves 3 dt just below the end)
        assert any(np.allclose(avg, np.mean(grid[:n], axis=0), rtol=0, atol=1e-15) for n in (4, 5))
        nominal = agent.get_action(time=0.0, averaging_duration=3 * dt, nominal_action=True)
        assert any(np.allclose(nominal, np.mean(grid[:n], axis=0), rtol=0, atol=1e-15) for n in (3, 4))
        assert np.array_equal(agent.get_action(time=0.0), a0)


def test_set_state_and_mocap(hostlib):
    with make("Particle") as agent:
        agent.set_state(time=0.3, qpos=[0.1, -0.05], qvel=[0.0, 0.2], mocap_pos=[0.2, 0.1, 0.01], mocap_quat=[1, 0, 0, 0])
        st = agent.get_state()
        assert st["time"] == 0.3 and list(st["qpos"]) == [0.1, -0.05] and list(st["mocap_pos"]) == [0.2, 0.1, 0.01]
        with pytest.raises(ValueError):
            agent.set_state(qpos=[1.0, 2.0, 3.0])
        agent.set_mocap({"goal": ([13, 14, 15], [1, 1, 1, 1])})
        st = agent.get_state()
        assert list(st["mocap_pos"]) == [13, 14, 15] and np.allclose(st["mocap_quat"], [0.5, 0.5, 0.5, 0.5])
        with pytest.raises(ValueError):
            agent.set_mocap({"pointmass": ([0, 0, 0], None)})


@pytest.mark.parametrize("name", ["Cross Entropy", "Robust Sampling"])
def test_other_planners_behind_the_same_agent(hostlib, name):
    with make("Cartpole", planner=name) as agent:
        agent.set_task_parameter("Goal", -0.5)
        s = agent.settings
        N, P = 10, min(s["num_spline_points"], 36)
        for _ in range(2):
            if name == "Cross Entropy":
                agent.planner.configure(num_trajectory=N, n_elite=3)
                agent.planner.SetNoise(agent._rng.standard_normal((N, P, agent.model.nu)))
            else:
                agent.planner.configure(ncandidates=2, nrepetitions=2)
                agent.planner.delegate.configure(num_trajectory=N)
                Pn = max(len(agent.planner.policy()[0]), 1)          # the robust planner never resamples: Reset()'s one node
                agent.planner.delegate.SetNoise(agent._rng.standard_normal((N, Pn, agent.model.nu)))
                agent.planner.SetForceNoise(agent._rng.standard_normal((4, s["steps"] - 1, 6 * agent.model.nbody)), s["steps"])
            agent.planner_step()
            agent.step()
        assert agent.get_state()["time"] > 0 and np.all(np.isfinite(agent.get_action()))


# ---- GPU: the same facade over the real engine, device noise (green on hardware since the round-1 driver run)
@pytest.mark.gpu
@pytest.mark.parametrize("planner_name", ["Sampling", "Cross Entropy", "Robust Sampling"])
def test_agent_on_the_device(planner_name):
    from tests.common import run_isolated
    run_isolated("tests.test_agent_api", "check_agent_on_the_device", planner_name, timeout=240)


def check_agent_on_the_device(planner_name):
    from mujoco_mpc_b200.agent import Agent
    with Agent("Particle", planner=planner_name) as agent:
        costs = []
        for _ in range(20):
            agent.planner_step()
            agent.step()
            costs.append(agent.get_total_cost())
        assert agent.get_state()["time"] == pytest.approx(20 * agent.plant_model.opt_timestep)
        assert np.all(np.isfinite(costs)) and np.all(np.isfinite(agent.get_action()))
        bt = agent.best_trajectory()
        assert bt["states"].shape == (51, 4) and bt["actions"].shape == (50, 2) and bt["times"].shape == (51,)
    with Agent("Cartpole") as agent:
        agent.set_task_parameter("Goal", -1.0)
        agent.set_cost_weights({"Vertical": 0, "Velocity": 0, "Centered": 0, "Control": 1})
        agent.planner_step()
        assert np.allclose(agent.get_action(), 0)


def test_modes(hostlib):
    # agent_test.py: test_get_set_default_mode / test_get_set_mode / test_get_all_modes / test_set_mode_error
    with make("Cartpole") as agent:
        agent.set_mode("default_mode")
        assert agent.get_mode() == "default_mode" and agent.get_all_modes() == ["default_mode"]
    with make("Quadruped Flat") as agent:
        assert tuple(agent.get_all_modes()) == ("Quadruped", "Biped", "Walk", "Scramble", "Flip")
        assert agent.get_mode() == "Quadruped"
        agent.set_mode("Walk")
        assert agent.get_mode() == "Walk"
        with pytest.raises(ValueError):
            agent.set_mode("Run")
        # the mode reaches the residual: the cost at the same state differs between modes
        c_walk = agent.get_total_cost()
        agent.set_mode("Biped")
        assert agent.get_total_cost() != c_walk
