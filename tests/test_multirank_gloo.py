"""World-size-2 test of the N>1 path on CPU (gloo): ranks own disjoint sample slices, exchange their
winner records with one all-gather and every rank picks the same global winner (lowest return,
ties -> lowest global sample index) -- the host logic bench.py and CudaSamplingPlanner::SetSharding
run over NCCL on the GPUs."""
import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from mujoco_mpc_b200 import planner
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100)                    # same stream on all ranks: the global table
    N, count = 64, 3 + 6
    out = []
    for trial in range(20):
        returns = rng.uniform(0, 10, world * N)
        if trial % 3 == 0:                              # exact ties across ranks
            returns[5] = returns[N + 7] = returns.min() - 1.0
        knots = rng.normal(0, 1, (world * N, count - 3))
        lo = rank * N
        local = returns[lo:lo + N]
        li = int(np.lexsort((np.arange(N), local))[0])
        mine = np.concatenate([[local[li], lo + li, local[0]], knots[lo + li]])
        t = torch.from_numpy(mine.copy()); allt = torch.empty(world * count, dtype=torch.float64)
        dist.all_gather_into_tensor(allt, t)
        rec = allt.numpy().reshape(world, count)
        best = planner.select_winner(rec)
        gi = int(rec[best, 1])
        expect = int(np.lexsort((np.arange(world * N), returns))[0])
        out.append((gi, expect, float(rec[0, 2]), float(returns[0]), bool(np.array_equal(rec[best, 3:], knots[expect]))))
    dist.destroy_process_group()
    q.put((rank, out))


def test_two_rank_minloc_gloo():
    import torch.multiprocessing as mp
    from mujoco_mpc_b200 import build
    build.build(); build.build_host()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    assert res[0] == res[1]                              # every rank reaches the same decision
    for gi, expect, nominal, nominal_expect, knots_ok in res[0]:
        assert gi == expect and nominal == nominal_expect and knots_ok


def _planner_worker(rank, world, port, q):
    """The sharded planning iteration itself: CudaSamplingPlanner (host class, compiled unchanged) on the host
    stand-in of the engine, SetSharding with a gloo all-gather; every rank injects its slice of ONE global noise
    tensor, so the result must equal a single-rank oracle iteration over world * N samples."""
    sys.path.insert(0, ROOT)
    import ctypes
    import torch
    import torch.distributed as dist
    from mujoco_mpc_b200 import planner, registry
    from tests.hostsim import driver
    from tests.common import synthetic_state
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    planner._LIB = planner.prototype(ctypes.CDLL(driver.build_planner_host()))
    model, task, s = registry.load("Cartpole")
    N, H, P, interp = 8, 31, min(s["num_spline_points"], 36), 2
    pl = planner.CudaSamplingPlanner(model, task)
    pl.configure(num_trajectory=N, interpolation=interp, num_spline_points=P, sigma0=s["exploration"], sigma1=0.0)
    pl.Reset(H, np.zeros(model.nu))

    def gather(mine):
        allt = torch.empty(world * mine.size, dtype=torch.float64)
        dist.all_gather_into_tensor(allt, torch.from_numpy(mine.copy()))
        return allt.numpy().reshape(world, -1)
    pl.SetSharding(rank, world, gather)
    rng = np.random.default_rng(9)                      # same stream on all ranks: the global noise tensor
    state, t, out = synthetic_state(model), 0.0, []
    for it in range(4):
        z = rng.standard_normal((world * N, P, model.nu))
        pl.SetNoise(z[rank * N:(rank + 1) * N]); pl.SetState(state, t); pl.OptimizePolicy(H)
        pt, pk = pl.policy()
        bt = pl.BestTrajectory()
        out.append((pl.winner, pl.improvement, pt.tolist(), pk.tolist(), None if bt is None else bt["total_return"]))
        t += 2 * s["timestep"]
    pl.close()
    dist.destroy_process_group()
    q.put((rank, out))


def test_two_rank_sharded_planner_equals_single_rank_oracle(oracle_lib):
    import torch.multiprocessing as mp
    from mujoco_mpc_b200 import registry
    from oracle import pyoracle
    from tests.common import synthetic_state
    from tests.hostsim import driver
    driver.build_planner_host()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    world = 2
    ps = [ctx.Process(target=_planner_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=300) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    # the single-rank oracle over the whole batch
    model, task, s = registry.load("Cartpole")
    N, H, P, interp = 8, 31, min(s["num_spline_points"], 36), 2
    o = pyoracle.Oracle(model, task); o.configure(P, interp, s["exploration"])
    o.set_policy(np.zeros(1), np.zeros((1, model.nu)), interp)
    rng = np.random.default_rng(9)
    state, t = synthetic_state(model), 0.0
    owners = set()
    for it in range(4):
        z = rng.standard_normal((world * N, P, model.nu))
        o.set_state(state, t); w = o.optimize(world * N, H, z)
        ret, _ = o.returns()
        ot, ok = o.policy(0)
        for r in range(world):
            winner, improvement, pt, pk, best_return = res[r][it]
            assert winner == w, (it, r)
            assert improvement == o.improvement()
            assert np.array_equal(np.array(pt), ot) and np.array_equal(np.array(pk), ok)
            # BestTrajectory() exists only on the rank that rolled the winner out
            if w // N == r:
                assert best_return == ret[w]; owners.add(r)
            else:
                assert best_return is None
        t += 2 * s["timestep"]
    assert owners                                          # at least one iteration's winner was checked
