#!/usr/bin/env python3
"""Algorithmic FP64 operations per sample-step, counted on the CPU oracle.

The oracle is rebuilt with every `double` replaced by a tallying wrapper (oracle/flopcount.h, `make -C oracle flops`); one
planning iteration under planning noise is rolled out single-threaded and the tally divided by samples x steps.  Writes
baseline/flops_<task>.json, which bench.py reads for the roofline numerator (`frac` = algorithmic; the kernel's executed
count from ncu stays beside it as `roofline_frac_executed`).

    python tools/count_flops.py [--samples 32]
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from oracle import pyoracle  # noqa: E402

TASKS = [("Cartpole", 101), ("Acrobot", 201), ("Particle", 51), ("Swimmer", 101), ("Walker", 101), ("Quadruped Flat", 101)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=32)
    a = ap.parse_args()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "flops"])
    L = ctypes.CDLL(os.path.join(ROOT, "oracle", "libmm_oracle_flops.so"))
    L.mmo_create.restype = ctypes.c_void_p; L.mmo_cost.restype = ctypes.c_double
    L.mmo_norm.restype = ctypes.c_double; L.mmo_improvement.restype = ctypes.c_double
    pyoracle._LIB = L                                    # the Oracle wrapper now drives the counting build
    os.makedirs(os.path.join(ROOT, "baseline"), exist_ok=True)
    for name, H in TASKS:
        model, task, s = registry.load(name)
        if name.startswith("Quadruped"):
            T.quadruped_snapshot(model, task, mode="Quadruped", gait="Trot", time=0.0)
        N, P = a.samples, min(s["num_spline_points"], 36)
        rng = np.random.default_rng(0)
        half = 0.5 * (model.actuator_ctrlrange[:, 1] - model.actuator_ctrlrange[:, 0])
        o = pyoracle.Oracle(model, task); o.configure(P, 2, s["exploration"])
        o.set_state(registry.synthetic_state(model), 0.0)
        o.set_policy(np.linspace(0, (H - 1) * s["timestep"], P), np.clip(rng.normal(0, 0.3, (P, model.nu)) * half, model.actuator_ctrlrange[:, 0], model.actuator_ctrlrange[:, 1]), 2)
        z = rng.standard_normal((N, P, model.nu))
        L.mmo_flops_reset()
        o.optimize(N, H, z, nthread=1)
        c = (ctypes.c_longlong * 5)(); L.mmo_flops(c)
        add, mul, div, sqrt, special = [int(x) for x in c]
        units = N * (H - 1)
        out = {"task": name, "samples": N, "horizon_steps": H, "sample_steps": units,
               "per_sample_step": {"add_sub": add / units, "mul": mul / units, "div": div / units, "sqrt": sqrt / units, "transcendental_equiv": special / units},
               "flops_per_sample_step": (add + mul + div + sqrt + special) / units,
               "how": "oracle rebuilt with a tallying double (oracle/flopcount.h); one planning iteration, exploration noise of the task, "
                      "steady-state random nominal; includes the terminal forward pass and the planner's own arithmetic; "
                      "transcendentals charged 20 (exp, log) / 30 (sin, cos) / 40 (pow, cosh, atan2, acos) operations"}
        path = os.path.join(ROOT, "baseline", "flops_" + name.lower().replace(" ", "_") + ".json")
        with open(path, "w") as f:
            json.dump(out, f, indent=1)
        print(f"{name:16s} {out['flops_per_sample_step']:12.1f} FP64 operations per sample-step   (add {add / units:.0f} mul {mul / units:.0f} div {div / units:.1f} sqrt {sqrt / units:.1f} special {special / units:.1f})")


if __name__ == "__main__":
    main()
