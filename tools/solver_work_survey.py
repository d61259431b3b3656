"""Where does the quadruped step spend its work?  Newton iterations and line-search evaluations per step (oracle, CPU).

    python tools/solver_work_survey.py [--task "Quadruped Flat"] [--rollouts 64] [--steps 100]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from oracle import pyoracle  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Quadruped Flat")
    ap.add_argument("--rollouts", type=int, default=64)
    ap.add_argument("--steps", type=int, default=100)
    a = ap.parse_args()
    pyoracle.build()
    m, t, s = registry.load(a.task)
    if a.task.startswith("Quadruped"):
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    m.v["opt_timestep"] = s["timestep"]; m._c = None
    o = pyoracle.Oracle(m, t)
    rng = np.random.default_rng(0)
    sigma = s["exploration"]
    lo, hi = m.actuator_ctrlrange[:, 0], m.actuator_ctrlrange[:, 1]
    half = 0.5 * (hi - lo)
    its, evs, nefc, ncon = [], [], [], []
    st = registry.synthetic_state(m)
    for r in range(a.rollouts):
        o.reset_data()
        o.set("qpos", st[:m.nq]); o.set("qvel", st[m.nq:m.nq + m.nv]); o.set("time", [0.0])
        for k in range(a.steps):
            if k % 10 == 0:
                o.set("ctrl", np.clip(sigma * half * rng.standard_normal(m.nu), lo, hi))
            e0 = o.get("dbg_line_evals")[0]
            o.step()
            its.append(int(o.get("solver_iter")[0])); evs.append(int(o.get("dbg_line_evals")[0] - e0))
            nefc.append(int(o.get("nefc")[0])); ncon.append(int(o.get("ncon")[0]))
    for name, v in (("newton iterations", its), ("line evaluations", evs), ("nefc", nefc), ("ncon", ncon)):
        v = np.array(v); qs = np.percentile(v, [50, 90, 99, 100])
        print(f"{name:18s} per step: mean {v.mean():.2f} median {qs[0]:.0f} p90 {qs[1]:.0f} p99 {qs[2]:.0f} max {qs[3]:.0f}")
    ev = np.array(evs); it = np.array(its)
    print(f"line evaluations per newton iteration: {ev.sum() / max(it.sum(), 1):.2f}")


if __name__ == "__main__":
    main()
