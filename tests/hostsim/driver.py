"""TEST INFRASTRUCTURE: ctypes driver for tests/hostsim/hostsim.cc (the engine's kernel-source
headers compiled with g++).  Used by the CPU test-suite and by hand to diff the device code path
against the oracle in the GPU-less container.  Not imported by the product."""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# PS_HOSTSIM_DEFINES="-DPS_ROLLOUT_SINGLE_SITE=1 ..." builds (and tests) an experimental configuration of the kernel
# source next to the default one
_EXTRA = os.environ.get("PS_HOSTSIM_DEFINES", "").split()
_TAG = ("_" + "".join(c if c.isalnum() else "_" for c in "".join(_EXTRA))) if _EXTRA else ""
_LIB = os.path.join(_HERE, f"libhostsim{_TAG}.so")
_CSRC = os.path.join(_HERE, "..", "..", "mujoco_mpc_b200", "csrc")


def build(force=False):
    src = os.path.join(_HERE, "hostsim.cc")
    deps = [src] + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    if not force and os.path.exists(_LIB) and all(os.path.getmtime(_LIB) >= os.path.getmtime(d) for d in deps):
        return _LIB
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared"] + _EXTRA + ["-x", "c++", src, "-o", _LIB]
    subprocess.run(cmd, check=True)
    return _LIB


def build_spec(force=False):
    """libhostsim_spec.so: the model-specialised lean instantiations under the defines of csrc/ps_variant_tiny_spec.cu"""
    src = os.path.join(_HERE, "hostsim_spec.cc"); out = os.path.join(_HERE, "libhostsim_spec.so")
    deps = [src] + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        return out
    subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++", src, "-o", out], check=True)
    return out


def rollout_spec(model, task, state, time, times, knots, interp, H, timestep, mocap=None):
    """One rollout on a model-specialised instantiation (raises when no specialisation matches the model / task)."""
    L = ctypes.CDLL(build_spec())
    ns, nu, nr = model.nq + model.nv + model.na, model.nu, task.num_residual
    state = np.ascontiguousarray(state, dtype=np.float64); times = np.ascontiguousarray(times, dtype=np.float64)
    knots = np.ascontiguousarray(knots, dtype=np.float64).reshape(len(times), nu)
    states = np.zeros((H, ns)); actions = np.zeros((H, nu)); residual = np.zeros((H, nr)); costs = np.zeros(H)
    total = ctypes.c_double(0); flags = ctypes.c_int(0); variant = ctypes.create_string_buffer(1024)
    mo = None if mocap is None else np.ascontiguousarray(mocap, dtype=np.float64)
    rc = L.hs_rollout_spec(ctypes.byref(model.c_struct()), ctypes.byref(task), _p(state), _p(mo) if mo is not None else None,
                           ctypes.c_double(time), _p(times), _p(knots), len(times), int(interp), int(H), ctypes.c_double(timestep),
                           _p(states), _p(actions), _p(residual), _p(costs), ctypes.byref(total), ctypes.byref(flags), variant, 1024)
    if rc != 0:
        raise RuntimeError(f"hs_rollout_spec failed ({rc}): {variant.value.decode()}")
    return dict(states=states, actions=actions, residual=residual, costs=costs, total=total.value, flags=flags.value,
                variant=variant.value.decode())


def build_planner_host(force=False):
    """libps_planner_host.so: the product's host planner classes (mujoco_mpc_b200/host/*.cc, unchanged) linked
    against hostengine.cc, the host-compiled stand-in of the engine ABI.  -Bsymbolic keeps its ps_* references
    inside the library even when the real libps_engine.so is loaded in the same process."""
    hdir = os.path.join(_HERE, "..", "..", "mujoco_mpc_b200", "host")
    out = os.path.join(_HERE, f"libps_planner_host{_TAG}.so")
    srcs = [os.path.join(_HERE, "hostengine.cc")] + [os.path.join(hdir, f) for f in sorted(os.listdir(hdir)) if f.endswith(".cc")]
    deps = srcs + [os.path.join(hdir, f) for f in os.listdir(hdir) if f.endswith(".h")] + \
        [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        return out
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wl,-Bsymbolic"] + _EXTRA + ["-o", out]
    for s in srcs:
        cmd += ["-x", "c++", s]
    subprocess.run(cmd + ["-lpthread"], check=True)
    return out


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def rollout(model, task, state, time, times, knots, interp, H, timestep, integrator, mocap=None, coop=False, xfrc=None):
    """One rollout of the device code path on the host.  knots: [P][nu].  Returns a dict with
    states [H][ns], actions [H][nu], residual [H][nr], costs [H], total, flags, variant."""
    L = ctypes.CDLL(build())
    ns, nu, nr = model.nq + model.nv + model.na, model.nu, task.num_residual
    state = np.ascontiguousarray(state, dtype=np.float64)
    times = np.ascontiguousarray(times, dtype=np.float64)
    knots = np.ascontiguousarray(knots, dtype=np.float64).reshape(len(times), nu)
    states = np.zeros((H, ns)); actions = np.zeros((H, nu)); residual = np.zeros((H, nr)); costs = np.zeros(H)
    total = ctypes.c_double(0); flags = ctypes.c_int(0); variant = ctypes.create_string_buffer(256)
    mo = None if mocap is None else np.ascontiguousarray(mocap, dtype=np.float64)
    fn = L.hs_rollout_coop if coop else L.hs_rollout
    zz = None
    if xfrc is not None:     # (z [H-1][6*nbody], std, rate): the force-perturbed instantiation of either coding
        zz = np.ascontiguousarray(xfrc[0], dtype=np.float64)
        L.hs_set_xfrc(_p(zz), ctypes.c_double(xfrc[1]), ctypes.c_double(xfrc[2]))
    else:
        L.hs_set_xfrc(None, ctypes.c_double(0), ctypes.c_double(1))
    rc = fn(ctypes.byref(model.c_struct()), ctypes.byref(task), _p(state), None if mo is None else _p(mo),
                      ctypes.c_double(time), _p(times), _p(knots), len(times), int(interp), int(H), ctypes.c_double(timestep),
                      int(integrator), _p(states), _p(actions), _p(residual), _p(costs), ctypes.byref(total),
                      ctypes.byref(flags), variant, 256)
    if rc != 0:
        raise RuntimeError(f"hs_rollout failed ({rc}): {variant.value.decode()}")
    return dict(states=states, actions=actions, residual=residual, costs=costs, total=total.value, flags=flags.value,
                variant=variant.value.decode())
