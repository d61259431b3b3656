"""Shared helpers for the parity tests: seeded synthetic inputs and the
oracle-side planning iteration with injected noise."""
import numpy as np
from mujoco_mpc_b200 import registry
from oracle import pyoracle


from mujoco_mpc_b200.registry import synthetic_state  # noqa: F401


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
