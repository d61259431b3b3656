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
