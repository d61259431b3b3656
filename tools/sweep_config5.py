"""BASELINE.json configs[4]: sample-count sweep x horizon sweep on cartpole and quadruped, one GPU.

    python tools/sweep_config5.py [--tasks Cartpole "Quadruped Flat"] [--cpu] > gpurun_out/sweep.md

Prints a markdown table: engine variant picked, rollout-kernel ms (CUDA events on the engine's
stream), whole-iteration wall ms through the C ABI (perturb + rollouts + argmin + winner D2H) and
sample-steps/s.  With --cpu the oracle planner (CPU restatement of the reference planner, all host
threads) is timed on a bounded sample (128 samples) of each horizon and printed beside it.
Noise comes from the on-device Philox stream; the nominal is a fixed random spline (steady state).
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from mujoco_mpc_b200.engine import Engine  # noqa: E402


def setup(name):
    m, t, s = registry.load(name)
    if name.startswith("Quadruped"):
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    return m, t, s


def nominal(m, s, H, P, interp):
    shift = max((H - 1) * s["timestep"] / (P - 1 if interp else P), 1e-5)
    times = shift * np.arange(P)
    half = 0.5 * (m.actuator_ctrlrange[:, 1] - m.actuator_ctrlrange[:, 0])
    knots = np.clip(np.random.default_rng(2024).normal(0, 0.3, (P, m.nu)) * half, m.actuator_ctrlrange[:, 0], m.actuator_ctrlrange[:, 1])
    return times, knots


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tasks", nargs="+", default=["Cartpole", "Quadruped Flat"])
    ap.add_argument("--samples", nargs="+", type=int, default=[128, 512, 2048, 8192, 32768, 65536, 131072, 262144])
    ap.add_argument("--horizons", nargs="+", type=int, default=[51, 101, 201])
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--budget-s", type=float, default=40.0, help="skip a cell whose predicted launch time exceeds this")
    a = ap.parse_args()
    print("| task | N | H | engine | rollout kernel ms | iteration ms (C ABI) | sample-steps/s | failures |")
    print("|---|---|---|---|---|---|---|---|")
    cpu_rows = []
    for name in a.tasks:
        m, t, s = setup(name)
        P, interp = min(s["num_spline_points"], 36), 2
        e = Engine(m, t)
        e.set_state(registry.synthetic_state(m), 0.0)
        e.set_seed(2024)
        rate = None
        for H in a.horizons:
            times, knots = nominal(m, s, H, P, interp)
            e.set_nominal(times, knots, interp)
            for N in a.samples:
                if rate and N * (H - 1) / rate > a.budget_s:
                    print(f"| {name} | {N} | {H} | (skipped: predicted {N * (H - 1) / rate:.0f} s) | | | | |")
                    continue
                reps = 3 if (not rate or N * (H - 1) / rate < 1.0) else 1
                best_ms, best_wall = 1e30, 1e30
                for _ in range(reps):
                    t0 = time.perf_counter()
                    e.rollouts(N, H, s["exploration"])
                    wall = (time.perf_counter() - t0) * 1e3
                    ms = e.kernel_ms()[1]
                    if ms < best_ms:
                        best_ms, best_wall = ms, wall
                _, fl = e.returns()
                r = N * (H - 1) / (best_wall * 1e-3)
                rate = max(rate or 0.0, r)
                print(f"| {name} | {N} | {H} | {e.variant()} | {best_ms:.3f} | {best_wall:.3f} | {r:.4g} | {int((fl & 1).sum())} |", flush=True)
        e.close()
        if a.cpu:
            from oracle import pyoracle
            pyoracle.build()
            cores = len(os.sched_getaffinity(0))
            for H in a.horizons:
                times, knots = nominal(m, s, H, P, interp)
                o = pyoracle.Oracle(m, t); o.configure(P, interp, s["exploration"])
                n_cpu = max(128, cores * 8)
                z = np.random.default_rng(1).standard_normal((n_cpu, P, m.nu))
                reps = 0; t0 = time.perf_counter()
                while time.perf_counter() - t0 < 4.0:
                    o.set_state(registry.synthetic_state(m), 0.0); o.set_policy(times, knots, interp); o.optimize(n_cpu, H, z, nthread=cores); reps += 1
                dt = time.perf_counter() - t0
                cpu_rows.append(f"| {name} | {n_cpu} | {H} | oracle planner, {cores} threads | | {dt / reps * 1e3:.3f} | {reps * n_cpu * (H - 1) / dt:.4g} | |")
    if cpu_rows:
        print("\nCPU arm (kind \"port\": reference planner logic over the restated MuJoCo step, bounded sample):\n")
        print("| task | N | H | engine | | iteration ms | sample-steps/s | |")
        print("|---|---|---|---|---|---|---|---|")
        print("\n".join(cpu_rows))


if __name__ == "__main__":
    main()
