"""How much constraint capacity do quadruped / walker rollouts actually use?

    python tools/contact_capacity_survey.py [--task "Quadruped Flat"] [--rollouts 256] [--steps 100]

Rolls the oracle (CPU) forward under spline-noise-like random controls from the synthetic planning state and from
perturbed states, and prints the distribution of active contacts (ncon), constraint rows (nefc) and contact rows
per step.  The capacities of the warp-per-rollout working set (Traits: MAXCON, MAXEFC, NJROW; csrc/ps_host_model.h)
decide how many rollouts fit an SM's shared memory, so this is the input for sizing a tighter capacity class.
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
    ap.add_argument("--rollouts", type=int, default=256)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--sigma", type=float, default=None, help="control noise as a fraction of the half range (default: the task's exploration)")
    a = ap.parse_args()
    pyoracle.build()
    m, t, s = registry.load(a.task)
    if a.task.startswith("Quadruped"):
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    o = pyoracle.Oracle(m, t)
    rng = np.random.default_rng(0)
    sigma = s["exploration"] if a.sigma is None else a.sigma
    lo, hi = m.actuator_ctrlrange[:, 0], m.actuator_ctrlrange[:, 1]
    half = 0.5 * (hi - lo)
    m.v["opt_timestep"] = s["timestep"]; m._c = None
    o = pyoracle.Oracle(m, t)
    ncon, nefc, fails = [], [], 0
    for r in range(a.rollouts):
        st = registry.synthetic_state(m, seed=r % 7, pos=0.1 * (r % 3), vel=0.3 * (r % 4))
        o.reset_data()
        o.set("qpos", st[:m.nq]); o.set("qvel", st[m.nq:m.nq + m.nv]); o.set("time", [0.0])
        base = np.clip(rng.normal(0, 0.3, m.nu) * half, lo, hi)
        for k in range(a.steps):
            if k % 10 == 0:
                ctrl = np.clip(base + sigma * half * rng.standard_normal(m.nu), lo, hi)
                o.set("ctrl", ctrl)
            o.step()
            q = np.array(o.get("qpos")[:m.nq])
            if not np.all(np.isfinite(q)) or np.abs(q).max() > 1e6:
                fails += 1; break
            ncon.append(int(o.get("ncon")[0])); nefc.append(int(o.get("nefc")[0]))
    ncon, nefc = np.array(ncon), np.array(nefc)
    print(f"{a.task}: {a.rollouts} rollouts x {a.steps} steps at dt {s['timestep']}, control noise {sigma} of the half range, {fails} diverged")
    for name, v in (("ncon", ncon), ("nefc", nefc)):
        qs = np.percentile(v, [50, 90, 99, 99.9, 100])
        print(f"  {name}: median {qs[0]:.0f}  p90 {qs[1]:.0f}  p99 {qs[2]:.0f}  p99.9 {qs[3]:.0f}  max {qs[4]:.0f}")


if __name__ == "__main__":
    main()
