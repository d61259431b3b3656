"""Would re-binning the quadruped rollouts by constraint count pay?  (oracle, CPU)

Rolls R rollouts of the headline shape (piecewise-constant random controls, H steps) on the oracle, records for each
(rollout, step) the solver's Newton iterations, line-search evaluations, nefc and ncon, and evaluates a lockstep cost model

    step cost of a group = max over its members of the per-step work proxy

for groups of 32 (a warp) and of `block` (a lockstep block) in natural order and after sorting by the previous step's key.

    python tools/binning_model.py [--rollouts 1024] [--steps 61] [--block 512]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from oracle import pyoracle  # noqa: E402


def record(task, R, H, seed=0):
    pyoracle.build()
    m, t, s = registry.load(task)
    if task.startswith("Quadruped"):
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    m.v["opt_timestep"] = s["timestep"]; m._c = None
    o = pyoracle.Oracle(m, t)
    rng = np.random.default_rng(seed)
    sigma = s["exploration"]
    lo, hi = m.actuator_ctrlrange[:, 0], m.actuator_ctrlrange[:, 1]
    half = 0.5 * (hi - lo)
    nominal = [np.clip(0.3 * half * rng.standard_normal(m.nu), lo, hi) for _ in range((H + 9) // 10 + 1)]
    st = registry.synthetic_state(m)
    out = np.zeros((4, R, H), dtype=np.int64)
    for r in range(R):
        o.reset_data()
        o.set("qpos", st[:m.nq]); o.set("qvel", st[m.nq:m.nq + m.nv]); o.set("time", [0.0])
        for k in range(H):
            if k % 10 == 0:
                o.set("ctrl", np.clip(nominal[k // 10] + sigma * half * rng.standard_normal(m.nu), lo, hi))
            e0 = o.get("dbg_line_evals")[0]
            o.step()
            out[0, r, k] = int(o.get("solver_iter")[0]); out[1, r, k] = int(o.get("dbg_line_evals")[0] - e0)
            out[2, r, k] = int(o.get("nefc")[0]); out[3, r, k] = int(o.get("ncon")[0])
    return out


def group_cost(w, order, g):
    """sum over groups of g consecutive entries of `order` of the group's max work, per step: returns the total"""
    R = len(order)
    ww = w[order]
    pad = (-R) % g
    if pad:
        ww = np.concatenate([ww, np.zeros(pad)])
    return ww.reshape(-1, g).max(axis=1).sum() * g


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Quadruped Flat")
    ap.add_argument("--rollouts", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=61)
    ap.add_argument("--block", type=int, default=512)
    a = ap.parse_args()
    rec = record(a.task, a.rollouts, a.steps)
    its, evs, nefc, ncon = rec
    R, H = its.shape
    # work proxies of one step: fixed part (kinematics, collision, smooth dynamics) + solver part
    proxies = {
        "newton iterations": its.astype(float),
        "line evaluations": evs.astype(float),
        "nefc": nefc.astype(float),
        "iters x nefc": (its * nefc).astype(float),
        "evals x nefc": (evs * nefc).astype(float),
    }
    for name, w in proxies.items():
        useful = w.sum()
        res = {}
        for label, g in (("warp", 32), ("block", a.block)):
            nat = sum(group_cost(w[:, k], np.arange(R), g) for k in range(H))
            srt = 0.0
            for k in range(H):
                key = nefc[:, k - 1] * 64 + its[:, k - 1] if k else np.zeros(R)
                srt += group_cost(w[:, k], np.argsort(key, kind="stable"), g)
            res[label] = (useful / max(nat, 1), useful / max(srt, 1))
        print(f"{name:18s} lane efficiency  warp: natural {res['warp'][0]:.2f} -> sorted {res['warp'][1]:.2f}   "
              f"block({a.block}): natural {res['block'][0]:.2f} -> sorted {res['block'][1]:.2f}")
    print(f"nefc mean {nefc.mean():.1f} max {nefc.max()}  iterations mean {its.mean():.2f} max {its.max()}  "
          f"steps with nefc == 0: {(nefc == 0).mean():.2f}")


if __name__ == "__main__":
    main()
