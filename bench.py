#!/usr/bin/env python
"""Benchmark of the Predictive Sampling planning iteration (BASELINE.json metric: rollout
sample-steps/s and planning-iteration latency).

A "step" is one full planning iteration (perturb -> N rollouts x H steps -> argmin) over every
task of the workload.  `value` is device-resident throughput (CUDA events on the engine's stream),
`e2e` goes through the host-buffer API (state + nominal H2D, winner record D2H inside the timed
region).  `--impl reference` times the CPU restatement of the reference planner (oracle/) on the
host cores.  One process per GPU under torchrun; samples are sharded across ranks (weak scaling)
and the winner is picked by an NCCL all-gather min-loc of the per-rank winner records.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1]: Acrobot + Particle, 4,096 samples, cubic spline, single B200
    "acrobot+particle_4096_cubic": [("Acrobot", 4096, 2, None), ("Particle", 4096, 2, None)],
    "cartpole_65536_1s": [("Cartpole", 65536, 2, 101)],
    "cartpole_10_zero": [("Cartpole", 10, 0, 101)],          # configs[0], the reference's CPU-runnable case
    "swimmer+walker_16384": [("Swimmer", 16384, 2, None), ("Walker", 16384, 2, None)],
    "quadruped_65536_1s": [("Quadruped Flat", 65536, 2, 101)],
}
DEFAULT_WORKLOAD = "acrobot+particle_4096_cubic"
METRIC = "rollout sample-steps/sec"
# FP64 flops per sample-step (FMA = 2), from the ncu instruction counts of the rollout kernel
# (profiles/flops_per_sample_step.json, see DESIGN.md "Measurement")
FLOPS_FILE = os.path.join(ROOT, "profiles", "flops_per_sample_step.json")


def load_flops():
    try:
        with open(FLOPS_FILE) as f:
            return json.load(f)
    except Exception:
        return {}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag or not self.rows:      # at least one sample, taken while the caller is still stepping
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_cases(workload, rank, world):
    from mujoco_mpc_b200 import registry
    from mujoco_mpc_b200.registry import synthetic_state
    cases = []
    for name, N, interp, H in WORKLOADS[workload]:
        model, task, s = registry.load(name)
        if name == "Quadruped Flat":   # planning state of a trotting robot heading for the goal
            from mujoco_mpc_b200 import task as _task
            _task.quadruped_snapshot(model, task, mode="Quadruped", gait="Trot", time=0.0)
        H = H or s["steps"]
        P = min(s["num_spline_points"], 36)
        state = synthetic_state(model)
        rng = np.random.default_rng(2024)
        # steady-state nominal: a previous winner, already resampled onto this iteration's knot times
        shift = max((H - 1) * s["timestep"] / (P - 1 if interp else P), 1e-5)
        times = 0.0 + shift * np.arange(P)
        half = 0.5 * (model.actuator_ctrlrange[:, 1] - model.actuator_ctrlrange[:, 0])
        knots = np.clip(rng.normal(0, 0.3, (P, model.nu)) * half, model.actuator_ctrlrange[:, 0], model.actuator_ctrlrange[:, 1])
        cases.append(dict(name=name, model=model, task=task, s=s, N=N, H=H, P=P, interp=interp, state=state, times=times,
                          knots=knots, sigma=s["exploration"], offset=rank * N))
    return cases


def run_reference(args, rank, world):
    """CPU arm: the reference planner restated over the restated MuJoCo step (oracle/), all host threads."""
    if rank != 0:
        return
    from oracle import pyoracle
    pyoracle.build()
    cases = make_cases(args.workload, 0, 1)
    cores = len(os.sched_getaffinity(0))
    # bounded sample: enough samples for ~1-2 s of CPU work per step
    runs = []
    for c in cases:
        o = pyoracle.Oracle(c["model"], c["task"])
        o.configure(c["P"], c["interp"], c["sigma"])
        n_cpu = min(c["N"], max(cores * 8, 128))
        z = np.random.default_rng(1).standard_normal((n_cpu, c["P"], c["model"].nu))
        runs.append((o, c, n_cpu, z))
    def step():
        for o, c, n_cpu, z in runs:
            o.set_state(c["state"], 0.0)
            o.set_policy(c["times"], c["knots"], c["interp"])
            o.optimize(n_cpu, c["H"], z, nthread=cores)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    units = sum(n * (c["H"] - 1) for _, c, n, _ in runs)
    v = units / dt
    sample = "; ".join(f"{c['name']} {n} of {c['N']} samples x {c['H'] - 1} steps" for _, c, n, _ in runs)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "sample-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": {"workload": args.workload},
            "cpu_baseline": {"value": v, "unit": "sample-steps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "sample-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    # BASELINE.json configs[4] (sample-count x horizon sweep, also under torchrun): one task at a chosen size
    ap.add_argument("--task", default=None, help='e.g. "Cartpole" or "Quadruped Flat": overrides --workload')
    ap.add_argument("--samples", type=int, default=65536, help="samples PER GPU (with --task)")
    ap.add_argument("--horizon", type=int, default=101, help="rollout steps incl. the terminal one (with --task)")
    ap.add_argument("--interpolation", type=int, default=2)
    args = ap.parse_args()
    if args.task:
        args.workload = f"{args.task.lower().replace(' ', '_')}_{args.samples}_{args.horizon}"
        WORKLOADS[args.workload] = [(args.task, args.samples, args.interpolation, args.horizon)]
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    from mujoco_mpc_b200.engine import Engine
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cases = make_cases(args.workload, rank, world)
    engines = []
    for c in cases:
        e = Engine(c["model"], c["task"], device=local)
        e.set_seed(2024)
        e.set_state(c["state"], 0.0); e.set_nominal(c["times"], c["knots"], c["interp"])
        engines.append(e)
    streams = [torch.cuda.ExternalStream(e.stream(), device=local) for e in engines]
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=f"cuda:{local}")   # 256 MB > 126 MB L2
    records = []
    for e, c in zip(engines, cases):
        ptr, n = e.winner_record_ptr()
        records.append((ptr, 3 + c["P"] * c["model"].nu))

    def record_tensor(k):
        # zero-copy view of the engine's device-resident winner record
        ptr, n = records[k]
        class _W:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}
        return torch.as_tensor(_W(), device=f"cuda:{local}")

    rec_views = [record_tensor(k) for k in range(len(cases))]
    gathered = [torch.empty(world * v.numel(), dtype=torch.float64, device=f"cuda:{local}") for v in rec_views]

    def device_step(timed):
        """one planning iteration per task; returns device ms (events on the engines' streams)"""
        ms = 0.0
        for k, (e, c) in enumerate(zip(engines, cases)):
            st = streams[k]
            with torch.cuda.stream(st):
                flush.zero_()                                     # L2 flush between iterations
                a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
                a.record(st)
                e.rollouts_async(c["N"], c["H"], c["sigma"], 0.0, c["offset"])
                if world > 1:
                    dist.all_gather_into_tensor(gathered[k], rec_views[k])   # min-loc over ranks (cost, index, knots)
                    g = gathered[k].view(world, -1)
                    tie = g[:, 0] == g[:, 0].min()                # lowest return, ties -> lowest global index
                    best = torch.where(tie, g[:, 1], torch.full_like(g[:, 1], float("inf"))).argmin()
                    _ = g[best]
                b.record(st)
            b.synchronize()
            ms += a.elapsed_time(b)
        return ms

    # end-to-end arm: the drop-in host planner (CudaSamplingPlanner), host buffers in and out
    from mujoco_mpc_b200.planner import CudaSamplingPlanner
    planners = []
    for c in cases:
        pl = CudaSamplingPlanner(c["model"], c["task"], device=local)
        pl.configure(num_trajectory=c["N"], interpolation=c["interp"], num_spline_points=c["P"], sigma0=c["sigma"], sigma1=0.0)
        pl.SetSeed(2024)
        if world > 1:
            cnt = 3 + c["P"] * c["model"].nu
            dev_in = torch.empty(cnt, dtype=torch.float64, device=f"cuda:{local}")
            dev_out = torch.empty(world * cnt, dtype=torch.float64, device=f"cuda:{local}")
            def gather(mine, dev_in=dev_in, dev_out=dev_out):
                dev_in.copy_(torch.from_numpy(mine)); dist.all_gather_into_tensor(dev_out, dev_in)
                return dev_out.cpu().numpy()
            pl.SetSharding(rank, world, gather)
        planners.append(pl)
    e2e_time = [0.0]

    def e2e_step():
        """public API: SetState (host state) -> OptimizePolicy (resample, H2D, K1-K3, D2H winner) -> ActionFromPolicy"""
        for pl, c in zip(planners, cases):
            pl.SetState(c["state"], e2e_time[0])
            pl.OptimizePolicy(c["H"])
            pl.ActionFromPolicy(e2e_time[0])
        e2e_time[0] += 0.01

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        device_step(False)
    launches0 = sum(e.launches() for e in engines)
    sampler = ClockSampler(local); sampler.start()
    barrier()
    ms_total = 0.0; kern_ms = np.zeros(3); lat = []
    for _ in range(args.steps):
        ms = device_step(True); ms_total += ms; lat.append(ms)
        for e in engines:
            kern_ms += np.array(e.kernel_ms())
    barrier()
    launches = sum(e.launches() for e in engines) - launches0
    t = torch.tensor([ms_total], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    units = sum(c["N"] * (c["H"] - 1) for c in cases) * world
    value = units / (ms_step * 1e-3)

    # end-to-end through the host-buffer API
    for _ in range(3):
        e2e_step()
    barrier(); t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier(); e2e_s = (time.perf_counter() - t0) / args.steps
    # the clocks sampler spans both timed regions (device-timed and end-to-end); if they were shorter than one
    # nvidia-smi query, keep stepping until it has seen the GPU under this load
    # (single rank only: under torchrun every device_step holds a collective, so the step count must be the
    # same on every rank -- a rank-local loop here deadlocks the all-gather)
    while world == 1 and len(sampler.rows) < 3 and time.perf_counter() - t0 < 5.0:
        device_step(False)
    clocks = sampler.summary()
    t = torch.tensor([e2e_s], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = units / float(t.item())
    # per iteration: the whole Constants block (model + task + plan, pinned) + nominal knots go up, the winner record comes back
    h2d = sum(e.constants_bytes() + 8 * c["P"] * c["model"].nu for e, c in zip(engines, cases))
    d2h = sum(8 * (3 + c["P"] * c["model"].nu) for c in cases)

    if rank == 0:
        flops = load_flops()
        peak = engines[0].fp64_peak()
        rollout_ms = kern_ms[1] / args.steps
        alg = sum(flops.get(c["name"], 0.0) * c["N"] * (c["H"] - 1) for c in cases)
        # DRAM bytes (read + write) of the rollout kernels of one step, from the ncu --set full captures
        dram = flops.get("_dram_bytes_per_sample_step", {})
        traffic = sum(dram.get(c["name"], float("nan")) * c["N"] * (c["H"] - 1) for c in cases)
        traffic = None if traffic != traffic else traffic
        achieved = alg / (rollout_ms * 1e-3) / 1e12 if alg else None
        # the same kernel against HBM: its measured DRAM traffic (mostly the thread-private working set of the larger
        # models spilling) over its duration, as a fraction of the measured copy bandwidth (MEASURED_PEAKS.json)
        hbm_peak = None
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                hbm_peak = float(json.load(f)["hbm_gbs"])
        except Exception:
            pass
        traffic_gbs = traffic / (rollout_ms * 1e-3) / 1e9 if (traffic and rollout_ms > 0) else None
        line = {"metric": METRIC, "value": value, "unit": "sample-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": {"workload": args.workload, "samples_per_gpu": [c["N"] for c in cases],
                                                "horizon_steps": [c["H"] for c in cases], "l2": "flushed (256 MB write) between iterations",
                                                "kernels": [e.variant() for e in engines],
                                                "parallelism": f"samples sharded over {world} rank(s), NCCL all-gather min-loc"},
                "planning_iteration_latency_ms": {"mean": ms_step, "p50": float(np.percentile(lat, 50)), "p95": float(np.percentile(lat, 95))}, "kernel_ms": {"perturb": kern_ms[0] / args.steps, "rollout": rollout_ms, "argmin": kern_ms[2] / args.steps},
                "e2e": {"value": e2e_value, "unit": "sample-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "latency_ms": e2e_s * 1e3},
                "gpu_launches": int(launches), "clocks": clocks,
                "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": (achieved / peak) if achieved else None,
                             "traffic": traffic, "traffic_unit": "bytes per step (sum of the step's rollout launches)", "peak_source": "DFMA probe measured in this run (MEASURED_PEAKS.json has no FP64 figure)",
                             "traffic_gbs": traffic_gbs, "hbm_peak_gbs": hbm_peak,
                             "traffic_frac_of_hbm": (traffic_gbs / hbm_peak) if (traffic_gbs and hbm_peak) else None}}
        if not args.no_cpu_baseline and world == 1:
            from oracle import pyoracle
            pyoracle.build()
            cores = len(os.sched_getaffinity(0))
            tot_units, tot_t, desc = 0, 0.0, []
            for c in cases:
                o = pyoracle.Oracle(c["model"], c["task"]); o.configure(c["P"], c["interp"], c["sigma"])
                n_cpu = min(c["N"], max(cores * 8, 128))
                z = np.random.default_rng(1).standard_normal((n_cpu, c["P"], c["model"].nu))
                reps = 0; t0 = time.perf_counter()
                while time.perf_counter() - t0 < 5.0:
                    o.set_state(c["state"], 0.0); o.set_policy(c["times"], c["knots"], c["interp"]); o.optimize(n_cpu, c["H"], z, nthread=cores); reps += 1
                tot_t += (time.perf_counter() - t0); tot_units += reps * n_cpu * (c["H"] - 1)
                desc.append(f"{c['name']}: {reps} iterations of {n_cpu} samples x {c['H'] - 1} steps")
            line["cpu_baseline"] = {"value": tot_units / tot_t, "unit": "sample-steps/s", "cores": cores, "kind": "port", "sample": "; ".join(desc)}
        print(json.dumps(line))
    for e in engines:
        e.close()
    for pl in planners:
        pl.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
