"""Shared helpers for the parity tests: seeded synthetic inputs and the
oracle-side planning iteration with injected noise."""
import numpy as np
from mujoco_mpc_b200 import registry
from oracle import pyoracle


def synthetic_state(model, seed=1234, pos=0.05, vel=0.1):
    rng = np.random.default_rng(seed)
    st = registry.home_state(model)
    st[:model.nq] += rng.uniform(-pos, pos, model.nq)
    st[model.nq:model.nq + model.nv] += rng.uniform(-vel, vel, model.nv)
    for j in range(model.njnt):        # renormalise free-joint quaternions
        if model.jnt_type[j] == 0:
            a = model.jnt_qposadr[j] + 3
            st[a:a + 4] /= np.linalg.norm(st[a:a + 4])
    return st


def oracle_iteration(model, task, settings, state, time, N, H, P, interp, sigma, z, nominal=None, mocap=None, nthread=1):
    """One OptimizePolicy on the oracle.  Returns (oracle, resampled nominal times, knots, winner)."""
    o = pyoracle.Oracle(model, task)
    o.configure(P, interp, sigma)
    o.set_state(state, time, mocap=mocap)
    if nominal is not None:
        o.set_policy(nominal[0], nominal[1], interp)
    w = o.optimize(N, H, z, nthread=nthread)
    nt, nk = o.policy(1)      # previous_policy = the resampled nominal of this iteration
    return o, nt, nk, w
