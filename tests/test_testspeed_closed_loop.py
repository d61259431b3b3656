"""BASELINE.json configs[0]: the reference's `testspeed` closed loop (mjpc/testspeed.cc:43-132) on
Cartpole -- 10 samples, 1 s horizon, zero-order spline -- with planner AND plant on the B200
(mujoco_mpc_b200/testspeed.py), against the same loop restated on the oracle under the same injected
noise: total cost, final state and every planned policy must agree."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


from tests.common import oracle_closed_loop  # noqa: E402


@pytest.mark.parametrize("task_name,interp,N,steps", [("Cartpole", 0, 10, 120), ("Acrobot", 2, 16, 120), ("Particle", 2, 16, 120),
                                                    ("Swimmer", 2, 16, 60), ("Walker", 2, 16, 60), ("Quadruped Flat", 2, 12, 40)])
def test_testspeed_closed_loop_matches_oracle(oracle_lib, task_name, interp, N, steps):
    from mujoco_mpc_b200.testspeed import synchronous_planning_cost
    cache = {}

    def noise(k, n, P, nu):
        if k not in cache:
            cache[k] = np.random.default_rng(1000 + k).standard_normal((n, P, nu))
        return cache[k]
    from mujoco_mpc_b200.testspeed import plant_model
    total_time = steps * plant_model(task_name).opt_timestep        # a plan after every 4th plant step
    g = synchronous_planning_cost(task_name, steps_per_planning_iteration=4, total_time=total_time, num_trajectory=N, interpolation=interp, noise=noise)
    o = oracle_closed_loop(task_name, 4, total_time, N, interp, noise)
    assert g["plans"] == o["plans"] == steps // 4
    assert abs(g["time"] - o["time"]) < 1e-12
    assert np.abs(g["state"] - o["state"]).max() <= 1e-9 * max(1.0, np.abs(o["state"]).max())
    assert abs(g["total_cost"] - o["total_cost"]) <= 1e-8 * max(1.0, abs(o["total_cost"]))


def test_testspeed_runs_with_device_noise():
    # the CLI path: on-device Philox noise, planner default settings; cost is finite and the loop plans every 4th step
    from mujoco_mpc_b200.testspeed import synchronous_planning_cost
    from mujoco_mpc_b200.testspeed import plant_model
    r = synchronous_planning_cost("Cartpole", steps_per_planning_iteration=4, total_time=50 * plant_model("Cartpole").opt_timestep, interpolation=0)
    assert r["plans"] == 13 and r["steps"] == 50 and np.isfinite(r["total_cost"]) and r["total_cost"] > 0
