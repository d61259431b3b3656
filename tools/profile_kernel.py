"""One rollout launch of one task, for ncu captures (B200_PROFILING.md recipe):

    python tools/profile_kernel.py --task "Quadruped Flat" --N 8192 --H 21            # plain run first
    ncu --set full --clock-control none --import-source on -k regex:rollout_kernel -c 1 \\
        -o gpurun_out/quad python tools/profile_kernel.py --task "Quadruped Flat" --N 8192 --H 21 --launches 1
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from mujoco_mpc_b200.engine import Engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Quadruped Flat")
    ap.add_argument("--N", type=int, default=8192)
    ap.add_argument("--H", type=int, default=21)
    ap.add_argument("--launches", type=int, default=3)
    ap.add_argument("--sigma", type=float, default=None)
    ap.add_argument("--nominal", choices=["zero", "random"], default="zero")
    a = ap.parse_args()
    m, t, s = registry.load(a.task)
    if a.task.startswith("Quadruped"):
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    P = min(s["num_spline_points"], 36)
    e = Engine(m, t)
    e.set_state(registry.synthetic_state(m), 0.0)
    times = np.linspace(0, (a.H - 1) * s["timestep"], P)
    half = 0.5 * (m.actuator_ctrlrange[:, 1] - m.actuator_ctrlrange[:, 0])
    knots = np.zeros((P, m.nu)) if a.nominal == "zero" else np.random.default_rng(0).normal(0, 0.3, (P, m.nu)) * half
    e.set_nominal(times, knots, s["interpolation"])
    e.set_seed(1)
    sigma = s["exploration"] if a.sigma is None else a.sigma
    for k in range(a.launches):
        t0 = time.perf_counter()
        w, wr, nr = e.rollouts(a.N, a.H, sigma)
        dt = time.perf_counter() - t0
        ms = e.kernel_ms()
        print(f"launch {k}: variant={e.variant()} winner={w} return={wr:.6g} kernel_ms={ms[1]:.3f} wall_ms={dt * 1e3:.3f} "
              f"sample-steps/s={a.N * (a.H - 1) / (ms[1] * 1e-3):.4g}")
    _, fl = e.returns()
    print("failures", int((fl & 1).sum()), "capacity-flagged", int(((fl & 2) != 0).sum()))
    e.close()


if __name__ == "__main__":
    main()
