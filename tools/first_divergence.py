"""Diagnostic: locate the first (sample, step) where the GPU trajectory departs from the oracle's bit pattern."""
import sys, numpy as np
sys.path.insert(0, '.')
from mujoco_mpc_b200 import registry
from mujoco_mpc_b200.engine import Engine
from oracle import pyoracle
from tests.common import synthetic_state, oracle_iteration
name = sys.argv[1] if len(sys.argv) > 1 else "Walker"
N = 256
model, task, s = registry.load(name)
H = s["steps"]; P = s["num_spline_points"]; sigma = s["exploration"]
z = np.random.default_rng(5).standard_normal((N, P, model.nu))
state = synthetic_state(model)
o, nt, nk, w = oracle_iteration(model, task, s, state, 0.37, N, H, P, 2, sigma, z, nthread=8)
e = Engine(model, task); e.set_state(state, 0.37); e.set_nominal(nt, nk, 2); e.set_noise(z)
e.rollouts(N, H, sigma, writeback=True)
tr = e.all_trajectories()
firsts = []
o2 = pyoracle.Oracle(model, task)
for i in range(N):
    ot = o.trajectory(i)
    d = np.abs(tr["states"][i] - ot["states"]).max(axis=1)
    nz = np.nonzero(d)[0]
    if len(nz) == 0:
        continue
    t = nz[0]
    firsts.append(t)
    if len(firsts) <= 6:
        # reproduce step t-1 -> t on the oracle and report contact state
        x = ot["states"][t - 1]
        o2.reset_data(); o2.set("qpos", x[:model.nq]); o2.set("qvel", x[model.nq:model.nq + model.nv])
        if model.na: o2.set("act", x[model.nq + model.nv:])
        o2.set("ctrl", ot["actions"][t - 1]); o2.set("time", ot["times"][t - 1])
        o2.forward()
        print(f"sample {i} first diff at step {t}: max|dstate|={d[t]:.3e} ncon={o2.get('ncon')[0]:.0f} nefc={o2.get('nefc')[0]:.0f} "
              f"iters={o2.get('solver_iter')[0]:.0f} which={np.nonzero(tr['states'][i][t] - ot['states'][t])[0]}  action diff {np.abs(tr['actions'][i][t-1]-ot['actions'][t-1]).max():.1e}"
              f" resid diff prev {np.abs(tr['residual'][i][t-1]-ot['residual'][t-1]).max():.1e}")
print("samples with any diff:", len(firsts), "of", N, " first-diff step histogram:", np.bincount(firsts)[:20] if firsts else None)
