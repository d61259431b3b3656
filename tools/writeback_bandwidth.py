"""Achieved HBM write bandwidth of the trajectory writeback (north_star evidence item): the rollout kernel in its
writeback instantiation stores states, actions, residuals, costs, times and traces of EVERY sample, SoA across
samples (coalesced: consecutive threads write consecutive addresses).

    python tools/writeback_bandwidth.py [--task Cartpole] [--samples 65536] [--horizon 101]

Prints the bytes one launch writes, the kernel time with and without writeback (CUDA events on the engine's stream)
and the bandwidth the difference corresponds to.  Run it under `ncu --set full -k regex:rollout_kernel` for the
dram__bytes_write.sum cross-check.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np  # noqa: E402
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from mujoco_mpc_b200.engine import Engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--task", default="Cartpole")
    ap.add_argument("--samples", type=int, default=65536)
    ap.add_argument("--horizon", type=int, default=101)
    a = ap.parse_args()
    m, t, s = registry.load(a.task)
    if a.task.startswith("Quadruped"):
        T.quadruped_snapshot(m, t, mode="Quadruped", gait="Trot", time=0.0)
    e = Engine(m, t)
    P, N, H = min(s["num_spline_points"], 36), a.samples, a.horizon
    e.set_state(registry.synthetic_state(m), 0.0)
    shift = max((H - 1) * s["timestep"] / (P - 1), 1e-5)
    e.set_nominal(shift * np.arange(P), np.zeros((P, m.nu)), 2)
    e.set_seed(1)
    ns = m.nq + m.nv + m.na
    per_step = ns + m.nu + t.num_residual + 1 + 1 + 3 * t.num_trace          # doubles per sample-step
    stride = (N + 31) & ~31
    total = 8.0 * per_step * H * stride
    out = {}
    for wb in (False, True):
        best = 1e30
        for _ in range(4):
            e.rollouts(N, H, s["exploration"], writeback=wb)
            best = min(best, e.kernel_ms()[1])
        out[wb] = best
    extra = out[True] - out[False]
    print(f"{a.task}: N={N} H={H} variant={e.variant()}  writeback record {per_step} doubles / sample-step, {total / 1e6:.1f} MB per launch")
    print(f"  rollout kernel {out[False]:.3f} ms cost-only, {out[True]:.3f} ms with writeback (+ the memset of the buffers is outside the kernel)")
    print(f"  writeback alone would need {total / 6.5e12 * 1e3:.3f} ms at the measured copy bandwidth (6.5 TB/s); it costs {extra:.3f} ms "
          f"= {total / max(extra * 1e-3, 1e-9) / 1e9:.0f} GB/s marginal, {total / (out[True] * 1e-3) / 1e9:.0f} GB/s averaged over the whole kernel")
    e.close()


if __name__ == "__main__":
    main()
