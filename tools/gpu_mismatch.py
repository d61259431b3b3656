"""Diagnostics: per-sample return mismatches between the CUDA engine and the oracle for one parity case."""
import argparse
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from mujoco_mpc_b200.engine import Engine  # noqa: E402
from tests.common import synthetic_state, oracle_iteration  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--task", default="Walker"); ap.add_argument("--N", type=int, default=1024); ap.add_argument("--H", type=int, default=0)
ap.add_argument("--repeat", type=int, default=2)
a = ap.parse_args()
model, task, s = registry.load(a.task)
if a.task.startswith("Quadruped"):
    T.quadruped_snapshot(model, task, mode="Quadruped", gait="Trot", time=0.37)
N, interp, seed, time = a.N, 2, 2024, 0.37
H = a.H or s["steps"]; P = min(s["num_spline_points"], 36); sigma = s["exploration"]
rng = np.random.default_rng(seed)
z = rng.standard_normal((N, P, model.nu))
state = synthetic_state(model)
nt = time - 0.013 + np.linspace(0, (H - 1) * s["timestep"], P)
nk = rng.uniform(-0.6, 0.6, (P, model.nu)) * 0.5 * (model.actuator_ctrlrange[:, 1] - model.actuator_ctrlrange[:, 0])
o, nt, nk, w = oracle_iteration(model, task, s, state, time, N, H, P, interp, sigma, z, nominal=(nt, nk), nthread=16)
oret, ofail = o.returns()
e = Engine(model, task)
print("variant", e.variant())
prev = None
for rep in range(a.repeat):
    e.set_state(state, time); e.set_nominal(nt, nk, interp); e.set_noise(z)
    gw, gwr, gnr = e.rollouts(N, H, sigma, writeback=(rep == a.repeat - 1))
    gret, gfail = e.returns()
    bad = np.nonzero(gret != oret)[0]
    print(f"rep {rep}: winner gpu {gw} oracle {w}; mismatching samples {len(bad)}; flagged {(gfail != 0).sum()}")
    for i in bad[:12]:
        print("   ", i, "gpu", gret[i], "oracle", oret[i], "flags", gfail[i], "ofail", ofail[i])
    if prev is not None:
        print("   run-to-run identical:", np.array_equal(prev, gret))
    prev = gret.copy()
if len(bad):
    i = int(bad[0]); tr = e.all_trajectories(); ot = o.trajectory(i)
    ds = np.abs(tr["states"][i] - ot["states"]).max(axis=1)
    first = int(np.argmax(ds > 0)) if (ds > 0).any() else -1
    print("sample", i, "first differing state step", first, "max diff", ds.max())
    dr = np.abs(tr["residual"][i] - ot["residual"]).max(axis=1)
    fr = int(np.argmax(dr > 0)) if (dr > 0).any() else -1
    print("   first differing residual step", fr)
    if fr >= 0:
        print("   residual gpu   ", tr["residual"][i][fr])
        print("   residual oracle", ot["residual"][fr])
        print("   residual diff  ", tr["residual"][i][fr] - ot["residual"][fr])
        print("   action diff at that step", np.abs(tr["actions"][i][fr] - ot["actions"][fr]).max(), "cost", tr["costs"][i][fr], ot["costs"][fr])
        da = np.abs(tr["actions"][i] - ot["actions"]).max(axis=1)
        print("   first differing action step", int(np.argmax(da > 0)) if (da > 0).any() else -1)
    for j in bad[1:6]:
        j = int(j); otj = o.trajectory(j)
        drj = np.abs(tr["residual"][j] - otj["residual"]); frj = int(np.argmax(drj.max(axis=1) > 0))
        dsj = np.abs(tr["states"][j] - otj["states"]).max(axis=1); fsj = int(np.argmax(dsj > 0)) if (dsj > 0).any() else -1
        print("   sample", j, "first residual diff step", frj, "components", np.nonzero(drj[frj] > 0)[0], "first state diff step", fsj)
e.close()
