#!/usr/bin/env python
"""Benchmark of the Predictive Sampling planning iteration (BASELINE.json metric: rollout
sample-steps/s and planning-iteration latency).

A "step" is one full planning iteration (perturb -> N rollouts x H steps -> argmin) over every task
of the workload; the tasks of a workload run concurrently on their engines' streams.  `value` is
device-resident throughput (CUDA events, L2 flushed between iterations), `e2e` goes through the
host-buffer API of the drop-in planner class (state + nominal H2D, winner record D2H inside the
timed region).  The default workload is the configuration north_star's target is quoted on:
65,536 samples x 1 s horizon on cartpole AND quadruped; per-task figures are in `tasks`.

One process per GPU under torchrun: the workload's samples are SHARDED across the ranks (strong
scaling: rank r rolls global samples [r N / R, (r + 1) N / R)), and the winner is picked by ONE
ncclAllGather of the per-rank winner records + a min-loc kernel inside the engine library
(ps_comm_init; lowest global index on ties, knots included).

`--impl reference` and the `cpu_baseline` leg time the CPU planner (oracle/: the reference planner
logic over the restated MuJoCo step; kind "port", physics unpinned) on all host cores, on the same
fraction of every task's samples so that the tasks weigh in as they do on the GPU.
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
    # the configuration north_star's target is quoted on: 65,536 samples x 1 s on cartpole AND quadruped
    # (under torchrun this is configs[3]: the quadruped's 65,536 samples sharded over the ranks)
    "cartpole+quadruped_65536_1s": [("Cartpole", 65536, 2, 101), ("Quadruped Flat", 65536, 2, 101)],
}
DEFAULT_WORKLOAD = "cartpole+quadruped_65536_1s"
METRIC = "rollout sample-steps/sec"
# FP64 operations per sample-step: `algorithmic` = counted on the CPU oracle by an instrumented double
# (baseline/flops_<task>.json), `executed` = the kernel's own dadd + dmul + 2 dfma from ncu
FLOPS_FILE = os.path.join(ROOT, "profiles", "flops_per_sample_step.json")
BASELINE_DIR = os.path.join(ROOT, "baseline")
TASK_NAMES = ["Cartpole", "Acrobot", "Particle", "Swimmer", "Walker", "Quadruped Flat"]


def load_flops():
    try:
        with open(FLOPS_FILE) as f:
            executed = json.load(f)
    except Exception:
        executed = {}
    algorithmic = {}
    for name in TASK_NAMES:
        p = os.path.join(BASELINE_DIR, "flops_" + name.lower().replace(" ", "_") + ".json")
        try:
            with open(p) as f:
                algorithmic[name] = float(json.load(f)["flops_per_sample_step"])
        except Exception:
            pass
    return executed, algorithmic


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


def make_cases(workload, rank, world, weak=False):
    """Per-rank view of the workload: rank r of R rolls global samples [offset, offset + N) of N_total."""
    from mujoco_mpc_b200 import registry
    from mujoco_mpc_b200.registry import synthetic_state
    cases = []
    for name, N_total, interp, H in WORKLOADS[workload]:
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
        if weak:
            N, offset, N_total = N_total, rank * N_total, N_total * world
        else:
            N = max(N_total // world, 1); offset = rank * N; N_total = N * world
        cases.append(dict(name=name, model=model, task=task, s=s, N=N, N_total=N_total, H=H, P=P, interp=interp, state=state,
                          times=times, knots=knots, sigma=s["exploration"], offset=offset))
    return cases


def _cpu_planners(cases, cores):
    """One CPU planner per task: the reference's OWN SamplingPlanner build (oracle/_ref/libmjpc_ref.so: mjpc/planners/sampling,
    trajectory.cc, task.cc, tasks/*.cc compiled unmodified; kind "reference") where it travelled with the snapshot, else the
    oracle's restatement of it (kind "port").  Both run the restated MuJoCo step below mj_step.  Returns (kind, [run(n)])."""
    from oracle import pyoracle
    pyoracle.build()
    L = pyoracle.ref_lib()
    have_ref = L is not None and hasattr(L, "refp_create") and os.environ.get("PS_CPU_ARM", "") != "port"
    runs = []
    for c in cases:
        if have_ref:
            from mujoco_mpc_b200 import registry, task as _task
            from tests.common import model_mocap
            r = pyoracle.RefPlanner(c["model"], registry._TASK_ID_NAME.get(c["name"], c["name"]))
            mocap = model_mocap(c["model"])
            if c["name"] == "Quadruped Flat":       # a trotting robot, as the GPU arm's snapshot
                par = r.task_spec()["parameters"].copy()
                par[_task.parameter_index(c["model"], "select_Gait")] = _task._as_double(_task.QUAD_GAITS["Trot"])
                r.task_set(mode=0, parameters=par)
                _, mocap = r.task_transition(c["state"], mocap, 0.0)
            cap = int(L.refp_max_trajectory())      # kMaxTrajectory: the reference rolls at most 128 samples per OptimizePolicy

            def run(n, r=r, c=c, mocap=mocap, cap=cap):
                done = 0
                while done < n:                     # a batch beyond kMaxTrajectory takes several planning iterations
                    k = min(cap, n - done)
                    r.configure(c["P"], c["interp"], c["sigma"], 0.0, k)
                    r.set_state(c["state"], 0.0, mocap=mocap); r.set_policy(c["times"], c["knots"], c["interp"])
                    r.optimize(k, c["H"], None, nthread=cores); done += k
            runs.append((run, r))
        else:
            o = pyoracle.Oracle(c["model"], c["task"]); o.configure(c["P"], c["interp"], c["sigma"])

            def run(n, o=o, c=c):
                z = np.random.default_rng(1).standard_normal((n, c["P"], c["model"].nu))
                o.set_state(c["state"], 0.0); o.set_policy(c["times"], c["knots"], c["interp"]); o.optimize(n, c["H"], z, nthread=cores)
            runs.append((run, o))
    return ("reference" if have_ref else "port"), runs


def cpu_arm(cases, cores, budget_s, steps=None, warmup=0):
    """The CPU planner on `cores` host threads over the SAME fraction f of every task's samples (so the tasks weigh in
    as on the GPU), f sized so that one step is about `budget_s` of wall time.  Returns (sample-steps/s, s per step,
    description, per-task rates, kind)."""
    kind, planners = _cpu_planners(cases, cores)
    if kind == "reference":
        # the reference ends the process (mju_error) when Ground() finds a foot more than 0.5 m under the floor
        # (utilities.cc:551, "SHOULD NOT OCCUR"); among 65,536 noisy rollouts a few do that.  Count and go on (the ray's -1 is
        # used as the distance, which is also what the oracle port and the device code do); the count is in `sample`.
        from oracle import pyoracle
        pyoracle.RefPlanner.tolerate_errors(True); pyoracle.RefPlanner.error_count(reset=True)
    runs = [[planners[i][0], c, 0, None, 0.0] for i, c in enumerate(cases)]
    # calibration: per-task rate on 8 rollouts per thread
    full_s = 0.0
    for r in runs:
        c = r[1]
        n0 = min(c["N_total"], max(8 * cores, 128))
        r[0](n0)
        t0 = time.perf_counter(); r[0](n0); dt = time.perf_counter() - t0
        full_s += c["N_total"] / n0 * dt
    f = min(1.0, budget_s / max(full_s, 1e-9))
    f = max(f, max(min(1.0, max(8 * cores, 128) / c["N_total"]) for c in cases))     # >= 8 rollouts per thread on every task
    for r in runs:
        c = r[1]
        r[2] = max(1, min(c["N_total"], int(round(f * c["N_total"]))))

    def step():
        for r in runs:
            t0 = time.perf_counter(); r[0](r[2]); r[4] += time.perf_counter() - t0
    for _ in range(warmup):
        step()
    for r in runs:
        r[4] = 0.0
    t0 = time.perf_counter(); n_steps = 0
    while (steps is not None and n_steps < steps) or (steps is None and (time.perf_counter() - t0 < budget_s * 6 or n_steps < 2)):
        step(); n_steps += 1
    dt = (time.perf_counter() - t0) / n_steps
    units = sum(r[2] * (r[1]["H"] - 1) for r in runs)
    per_task = {r[1]["name"]: r[2] * (r[1]["H"] - 1) * n_steps / r[4] for r in runs}
    how = "the reference's own SamplingPlanner + ThreadPool, kMaxTrajectory = 128 samples per OptimizePolicy" if kind == "reference" else \
        "persistent thread pool, dynamic scheduling"
    sample = f"{n_steps} steps of " + "; ".join(f"{r[1]['name']} {r[2]} of {r[1]['N_total']} samples x {r[1]['H'] - 1} steps" for r in runs) + \
        f" (fraction {f:.4g} of every task, {cores} threads, {how})"
    if kind == "reference":
        from oracle import pyoracle
        nerr = pyoracle.RefPlanner.error_count(reset=True); pyoracle.RefPlanner.tolerate_errors(False)
        if nerr:
            sample += (f"; {nerr} mju_error call(s) raised inside the reference's code (Ground(): no geom under a foot, utilities.cc:551) "
                       "were counted and passed over instead of ending the process")
    return units / dt, dt, sample, per_task, kind


CPU_NOTE = {"reference": "the reference's own planner / rollout / cost / residual code (compiled unmodified into oracle/_ref) over the restated "
                         "MuJoCo 3.1.4 step: MuJoCo itself is absent, so the physics below mj_step is the oracle's (unpinned)",
            "port": "reference planner logic restated over the restated MuJoCo 3.1.4 step (physics unpinned); not the reference binary"}


def run_reference(args, rank, world):
    """CPU arm: the reference planner restated over the restated MuJoCo step (oracle/), all host threads."""
    if rank != 0:
        return
    cases = make_cases(args.workload, 0, 1)
    cores = len(os.sched_getaffinity(0))
    v, dt, sample, per_task, kind = cpu_arm(cases, cores, budget_s=1.5, steps=args.steps, warmup=args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "sample-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": {"workload": args.workload},
            "tasks": [{"name": k, "value": x, "unit": "sample-steps/s"} for k, x in per_task.items()],
            "cpu_baseline": {"value": v, "unit": "sample-steps/s", "cores": cores, "kind": kind, "sample": sample, "note": CPU_NOTE[kind]},
            "e2e": {"value": v, "unit": "sample-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--weak", action="store_true", help="every rank rolls the workload's full sample count (weak scaling)")
    # BASELINE.json configs[4] (sample-count x horizon sweep, also under torchrun): one task at a chosen size
    ap.add_argument("--task", default=None, help='e.g. "Cartpole" or "Quadruped Flat": overrides --workload')
    ap.add_argument("--samples", type=int, default=65536, help="total samples (with --task); per GPU with --weak")
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
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    cases = make_cases(args.workload, rank, world, weak=args.weak)
    engines = []
    for c in cases:
        e = Engine(c["model"], c["task"], device=local)
        e.set_seed(2024)
        e.set_state(c["state"], 0.0); e.set_nominal(c["times"], c["knots"], c["interp"])
        if world > 1:
            e.comm_init_torch(dist, rank, world)          # the library's own NCCL communicator for the winner exchange
        engines.append(e)
    streams = [torch.cuda.ExternalStream(e.stream(), device=local) for e in engines]
    main_stream = torch.cuda.current_stream(dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)   # 256 MB > 126 MB L2

    def device_step():
        """one planning iteration of every task, the tasks concurrently on their engines' streams; returns device ms"""
        flush.zero_()                                             # L2 flush between iterations
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(main_stream)
        for k, (e, c) in enumerate(zip(engines, cases)):
            streams[k].wait_event(a)
            e.rollouts_async(c["N"], c["H"], c["sigma"], 0.0, c["offset"])   # K1-K3 (+ the all-gather min-loc when sharded)
            done = torch.cuda.Event(); done.record(streams[k]); main_stream.wait_event(done)
        b.record(main_stream)
        b.synchronize()
        return a.elapsed_time(b)

    # end-to-end arm: the drop-in host planner (CudaSamplingPlanner), host buffers in and out
    from mujoco_mpc_b200.planner import CudaSamplingPlanner
    from mujoco_mpc_b200 import engine as _eng
    elib = _eng.lib()
    planners = []
    for c in cases:
        pl = CudaSamplingPlanner(c["model"], c["task"], device=local)
        pl.configure(num_trajectory=c["N"], interpolation=c["interp"], num_spline_points=c["P"], sigma0=c["sigma"], sigma1=0.0)
        pl.SetSeed(2024)
        if world > 1:
            pl.SetShardingNative(rank, world, dist)       # ncclAllGather inside the engine, no host bounce
        planners.append(pl)
    e2e_time = [0.0]
    e2e_dev_ms = [0.0]                 # device time of the K1 + K2 + K3 launches inside the end-to-end steps

    def e2e_step():
        """public API: SetState (host state) -> OptimizePolicy (resample, H2D, K1-K3, D2H winner) -> ActionFromPolicy"""
        import ctypes
        for pl, c in zip(planners, cases):
            pl.SetState(c["state"], e2e_time[0])
            pl.OptimizePolicy(c["H"])
            pl.ActionFromPolicy(e2e_time[0])
            a, b, d = ctypes.c_float(), ctypes.c_float(), ctypes.c_float()
            if elib.ps_last_kernel_ms(ctypes.c_void_p(pl.engine_ptr()), ctypes.byref(a), ctypes.byref(b), ctypes.byref(d)) == 0:
                e2e_dev_ms[0] += a.value + b.value + d.value
        e2e_time[0] += 0.01

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        device_step()
    launches0 = sum(e.launches() for e in engines)
    sampler = ClockSampler(local); sampler.start()
    barrier()
    ms_total = 0.0; kern_ms = np.zeros((len(cases), 4)); lat = []
    for _ in range(args.steps):
        ms = device_step(); ms_total += ms; lat.append(ms)
        for k, e in enumerate(engines):
            kern_ms[k] += np.array(e.kernel_ms4())
    barrier()
    launches = sum(e.launches() for e in engines) - launches0
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    units = sum(c["N_total"] * (c["H"] - 1) for c in cases)
    value = units / (ms_step * 1e-3)
    kern_ms /= args.steps
    kt = torch.tensor(kern_ms, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(kt, op=dist.ReduceOp.MAX)
    kern_ms = kt.cpu().numpy()

    # end-to-end through the host-buffer API
    for _ in range(3):
        e2e_step()
    e2e_dev_ms[0] = 0.0
    barrier(); t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier(); e2e_s = (time.perf_counter() - t0) / args.steps
    e2e_device_ms = e2e_dev_ms[0] / args.steps
    # the clocks sampler spans both timed regions (device-timed and end-to-end); if they were shorter than ~20
    # nvidia-smi queries, keep stepping until it has seen the GPU under this load (every rank runs the same number
    # of extra steps: each sharded step holds a collective)
    extra = 0
    if len(sampler.rows) < 20:
        extra = int(min(500, 3.0 / max(ms_step * 1e-3, 1e-4)))
    if world > 1:
        et = torch.tensor([extra], dtype=torch.int64, device=dev); dist.all_reduce(et, op=dist.ReduceOp.MAX); extra = int(et.item())
    for _ in range(extra):
        device_step()
    clocks = sampler.summary()
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = units / float(t.item())
    # per iteration: the plan part of the Constants block (state + knot times, pinned) + nominal knots go up, the winner record comes back
    h2d = sum(e.upload_bytes() + 8 * c["P"] * c["model"].nu for e, c in zip(engines, cases))
    d2h = sum(8 * (3 + c["P"] * c["model"].nu) for c in cases)

    if rank == 0:
        executed, algorithmic = load_flops()
        peak = engines[0].fp64_peak()
        hbm_peak = None
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                hbm_peak = float(json.load(f)["hbm_gbs"])
        except Exception:
            pass
        dram = executed.get("_dram_bytes_per_sample_step", {})
        tasks = []
        for k, (e, c) in enumerate(zip(engines, cases)):
            su = c["N_total"] * (c["H"] - 1)
            rollout_s = kern_ms[k][1] * 1e-3
            alg = algorithmic.get(c["name"]); exe = executed.get(c["name"])
            tasks.append({"name": c["name"], "samples": c["N_total"], "samples_per_gpu": c["N"], "horizon_steps": c["H"], "kernel": e.variant(),
                          "value": su / (sum(kern_ms[k]) * 1e-3), "unit": "sample-steps/s",
                          "kernel_ms": {"perturb": kern_ms[k][0], "rollout": kern_ms[k][1], "argmin": kern_ms[k][2], "exchange": kern_ms[k][3]},
                          "flops_per_sample_step": {"algorithmic": alg, "executed": exe},
                          "roofline_frac": (alg * su / rollout_s / 1e12 / peak / world) if alg else None,
                          "roofline_frac_executed": (exe * su / rollout_s / 1e12 / peak / world) if exe else None,
                          "dram_bytes_per_sample_step": dram.get(c["name"])})
        # the dominant kernel = the rollout kernel of the task that takes the longest
        dom = int(np.argmax(kern_ms[:, 1])); dc = cases[dom]; dsu = dc["N_total"] * (dc["H"] - 1)
        dflops = algorithmic.get(dc["name"]) or executed.get(dc["name"])
        achieved = (dflops * dsu / world / (kern_ms[dom][1] * 1e-3) / 1e12) if dflops else None
        traffic = dram.get(dc["name"]) * dsu / world if dram.get(dc["name"]) is not None else None
        line = {"metric": METRIC, "value": value, "unit": "sample-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak" if args.weak else "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": {"workload": args.workload, "samples": [c["N_total"] for c in cases],
                                                "samples_per_gpu": [c["N"] for c in cases],
                                                "horizon_steps": [c["H"] for c in cases], "l2": "flushed (256 MB write) between iterations",
                                                "kernels": [e.variant() for e in engines],
                                                "parallelism": f"samples sharded over {world} rank(s); ncclAllGather min-loc of the winner records inside the engine library"},
                "tasks": tasks,
                "planning_iteration_latency_ms": {"mean": ms_step, "p50": float(np.percentile(lat, 50)), "p95": float(np.percentile(lat, 95))},
                "e2e": {"value": e2e_value, "unit": "sample-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "latency_ms": e2e_s * 1e3, "device_ms": e2e_device_ms, "host_overhead_ms": e2e_s * 1e3 - e2e_device_ms,
                        "note": "closed loop: every step re-plans from the previous winner, so its rollouts (and the contact load of the quadruped's) "
                                "differ from the fixed nominal of the device-timed steps; device_ms is the K1+K2+K3 time of THESE steps, "
                                "host_overhead_ms what the host-buffer API adds on top"},
                "gpu_launches": int(launches), "clocks": clocks,
                "roofline": {"bound": "fp64", "kernel": f"{engines[dom].variant()} rollout kernel ({dc['name']})", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                             "frac": (achieved / peak) if achieved else None,
                             "flops": "algorithmic (oracle-counted, baseline/flops_*.json)" if algorithmic.get(dc["name"]) else "executed (ncu)",
                             "traffic": traffic, "traffic_unit": "DRAM bytes per launch (ncu --set full, profiles/)",
                             "peak_source": "DFMA probe measured in this run (MEASURED_PEAKS.json has no FP64 figure); with -fmad=false (the parity policy) the attainable peak is half of it",
                             "hbm_peak_gbs": hbm_peak,
                             "traffic_gbs": (traffic / (kern_ms[dom][1] * 1e-3) / 1e9) if traffic else None}}
        if not args.no_cpu_baseline and world == 1:
            cores = len(os.sched_getaffinity(0))
            v, dt, sample, per_task, kind = cpu_arm(make_cases(args.workload, 0, 1), cores, budget_s=2.0)
            line["cpu_baseline"] = {"value": v, "unit": "sample-steps/s", "cores": cores, "kind": kind, "sample": sample,
                                    "per_task": per_task, "note": CPU_NOTE[kind]}
            if kind == "reference":      # the restated planner (one batch, dynamic scheduling, no 128-sample cap) beside it, for scale
                os.environ["PS_CPU_ARM"] = "port"
                v2, _, sample2, per_task2, _ = cpu_arm(make_cases(args.workload, 0, 1), cores, budget_s=1.0)
                os.environ.pop("PS_CPU_ARM", None)
                line["cpu_baseline"]["port"] = {"value": v2, "per_task": per_task2, "sample": sample2, "note": CPU_NOTE["port"]}
        print(json.dumps(line))
    for e in engines:
        e.close()
    for pl in planners:
        pl.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
