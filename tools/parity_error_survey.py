import sys, numpy as np
sys.path.insert(0,'/root/repo')
from mujoco_mpc_b200 import registry
from mujoco_mpc_b200.engine import Engine
from tests.common import synthetic_state, oracle_iteration
for name, N in [("Walker", 4096), ("Swimmer", 1024), ("Cartpole", 2048)]:
    model, task, s = registry.load(name)
    H = s["steps"]; P = s["num_spline_points"]; sigma = s["exploration"]
    rng = np.random.default_rng(5)
    z = rng.standard_normal((N, P, model.nu))
    state = synthetic_state(model)
    o, nt, nk, w = oracle_iteration(model, task, s, state, 0.37, N, H, P, 2, sigma, z, nthread=16)
    e = Engine(model, task); e.set_state(state, 0.37); e.set_nominal(nt, nk, 2); e.set_noise(z)
    gw, gwr, gnr = e.rollouts(N, H, sigma, writeback=True)
    oret, ofail = o.returns(); gret, gfail = e.returns()
    err = np.abs(gret - oret) / np.maximum(1, np.abs(oret))
    print(name, 'winner', gw, w, 'fail', ofail.sum(), (gfail&1).sum(), 'cost relerr max %.2e  p99 %.2e median %.2e  n>1e-12: %d  exact: %d' % (err.max(), np.quantile(err, .99), np.median(err), (err > 1e-12).sum(), (err == 0).sum()))
    e.close()
