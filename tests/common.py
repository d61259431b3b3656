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


def model_mocap(model):
    """mocap pose [nmocap][7] of the model as loaded (what mjData holds after mj_resetData)"""
    out = np.zeros((max(model.nmocap, 0), 7))
    for b in range(model.nbody):
        i = int(model.body_mocapid[b])
        if i >= 0:
            out[i, :3] = np.asarray(model.body_pos).reshape(-1, 3)[b]
            out[i, 3:] = np.asarray(model.body_quat).reshape(-1, 4)[b]
    return out.reshape(-1) if model.nmocap else None


def run_isolated(module, func, *args, timeout=900):
    """Run `module.func(*args)` in a child interpreter.  Used by GPU tests whose CUDA path has not run on hardware
    yet: a device fault or an abort() in there must not take the verified tests of the same session down with it."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = f"import {module} as m; m.{func}(*{args!r}); print('ISOLATED-OK')"
    r = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0 and "ISOLATED-OK" in r.stdout, f"exit {r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"
