#!/usr/bin/env python3
"""Generate tests/golden/planner_reference.npz from the reference's OWN Predictive Sampling code: mjpc::SamplingPlanner
(OptimizePolicy = UpdateNominalPolicy + AddNoiseToPolicy + Rollouts + argmin), SamplingPolicy / TimeSpline, Trajectory::
NoisyRollout + UpdateReturn, Task / BaseResidualFn cost code, State and the six task residuals (incl. the quadruped's
Transition-managed modes), compiled UNMODIFIED from /root/reference into oracle/_ref/libmjpc_ref.so (oracle/Makefile: ref).
Below mj_step the physics is the oracle's restated MuJoCo step -- the fixtures pin the planner / rollout / cost / residual
side of the oracle and of the CUDA engine, not the physics.

Every case: seeded synthetic state, injected standard-normal knot noise (replayed through the absl shim on a one-thread
pool), two consecutive planning iterations (the second exercises the resampling of the previous plan).  Stored per
iteration: all returns + failure flags, the winner, policy and previous policy after the iteration, and the full
trajectories (states / actions / residual / costs / trace) of samples 0, 1, the winner and N-1.

Run in the build container (where /root/reference exists); the committed .npz travels to the GPU box.
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mujoco_mpc_b200 import registry, task as T  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.common import model_mocap  # noqa: E402

# (case id, task, interpolation, N, H, sigma2, quadruped (mode, gait) or None)
CASES = [("cartpole_cubic", "Cartpole", 2, 24, 51, 0.0, None), ("cartpole_zero", "Cartpole", 0, 10, 101, 0.0, None),
         ("cartpole_linear_std2", "Cartpole", 1, 24, 41, 0.9, None), ("acrobot_cubic", "Acrobot", 2, 24, 61, 0.0, None),
         ("particle_cubic", "Particle", 2, 24, 51, 0.0, None), ("swimmer_cubic", "Swimmer", 2, 16, 41, 0.0, None),
         ("walker_cubic", "Walker", 2, 16, 51, 0.0, None), ("quadruped_trot", "Quadruped Flat", 2, 12, 41, 0.0, ("Quadruped", "Trot")),
         ("quadruped_biped", "Quadruped Flat", 2, 8, 31, 0.0, ("Biped", "Trot")), ("quadruped_walk_canter", "Quadruped Flat", 2, 8, 31, 0.0, ("Walk", "Canter")),
         ("quadruped_scramble_gallop", "Quadruped Flat", 2, 8, 31, 0.0, ("Scramble", "Gallop")), ("quadruped_flip", "Quadruped Flat", 2, 8, 31, 0.0, ("Flip", "Trot"))]


def reference_case(cid, name, interp, N, H, sigma2, quad, seed):
    """Runs one case on the reference build.  Returns (inputs, outputs) dicts of arrays."""
    model, task, s = registry.load(name)
    P = min(s["num_spline_points"], 36)
    rng = np.random.default_rng(seed)
    state = registry.synthetic_state(model, seed=seed)
    sigma = s["exploration"]; time = 0.37
    r = pyoracle.RefPlanner(model, registry._TASK_ID_NAME.get(name, name))
    mocap = model_mocap(model)
    if quad is not None:
        mode, gait = quad
        par = r.task_spec()["parameters"].copy()
        par[T.parameter_index(model, "select_Gait")] = T._as_double(T.QUAD_GAITS[gait])
        r.task_set(mode=0, parameters=par)
        _, mocap = r.task_transition(state, mocap, time - 0.01)          # the first Transition only accepts Quadruped / Biped
        r.task_set(mode=T.QUAD_MODES[mode])
        md, mocap = r.task_transition(state, mocap, time)
        assert md == T.QUAD_MODES[mode], (cid, md)
    spec = r.task_spec()
    inp = dict(task=name, interp=interp, N=N, H=H, P=P, sigma=sigma, sigma2=sigma2, state=state, time0=time, dt_iter=0.02,
               mocap=np.zeros(0) if mocap is None else mocap, parameters=spec["parameters"], weight=spec["weight"],
               aux=r.quadruped_state() if quad is not None else np.zeros(0))
    r.configure(P, interp, sigma, sigma2, N)
    out = {}
    for it in range(2):
        z = rng.standard_normal((N, P, model.nu)); std2 = (rng.uniform(size=N) < 0.2).astype(np.uint8)
        inp[f"z{it}"] = z; inp[f"std2_{it}"] = std2
        t = time + it * 0.02
        r.set_state(state, t, mocap=mocap)
        w = r.optimize(N, H, z, use_std2=std2 if sigma2 > 0 else None)
        ret, fail = r.returns()
        out[f"returns{it}"] = ret; out[f"failure{it}"] = fail; out[f"winner{it}"] = np.int64(w); out[f"improvement{it}"] = r.improvement()
        pt, pk = r.policy(0); qt, qk = r.policy(1)
        out[f"policy_times{it}"] = pt; out[f"policy_knots{it}"] = pk; out[f"previous_times{it}"] = qt; out[f"previous_knots{it}"] = qk
        keep = sorted(set([0, 1, w, N - 1]))
        out[f"kept{it}"] = np.array(keep)
        for i in keep:
            tr = r.trajectory(i)
            for k in ("states", "actions", "residual", "costs", "trace", "times"):
                out[f"traj{it}_{i}_{k}"] = tr[k]
    return inp, out


def main():
    pyoracle.build()
    pyoracle.RefPlanner.zero_warmstart_per_rollout(True)      # the engine's rollout semantics (deviation 2)
    blob = {"cases": np.array([c[0] for c in CASES])}
    for k, (cid, name, interp, N, H, sigma2, quad) in enumerate(CASES):
        inp, out = reference_case(cid, name, interp, N, H, sigma2, quad, seed=100 + k)
        for key, v in inp.items():
            blob[f"{cid}/in/{key}"] = np.asarray(v)
        for key, v in out.items():
            blob[f"{cid}/out/{key}"] = np.asarray(v)
        print(cid, "winners", int(out["winner0"]), int(out["winner1"]), "best return", float(out["returns1"].min()))
    path = os.path.join(ROOT, "tests", "golden", "planner_reference.npz")
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
