"""ps_argmin_noalloc (include/ps_engine.h): the allocation-free winner selection whose scratch is a file-scope
__device__ array.  No counterpart in the reference.  Checked through
the C ABI against the existing entry point (ps_rollouts' K2 + K3 winner) on two calls in a row; collected last so that
a failure here cannot hide any other test under `pytest -x`."""
import numpy as np
import pytest


@pytest.mark.gpu
def test_noalloc_argmin_matches_engine_winner_twice():
    import torch
    from mujoco_mpc_b200 import registry
    from mujoco_mpc_b200 import engine as eng
    from mujoco_mpc_b200.engine import Engine
    model, task, s = registry.load("Cartpole")
    N, H, P = 5000, 31, 6                       # not a multiple of the 1024 returns one block reduces
    e = Engine(model, task, device=0)
    e.set_state(registry.synthetic_state(model), 0.0)
    e.set_nominal(np.linspace(0.0, 0.3, P), np.zeros((P, model.nu)), 2)
    dev = torch.device("cuda", 0)
    out = torch.zeros(2, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    for call, offset in enumerate((0, 7000)):
        e.set_noise(np.random.default_rng(100 + call).standard_normal((N, P, model.nu)))
        w, wr, _ = e.rollouts(N, H, 0.4, sample_offset=offset)          # existing entry point: K2 block min-loc + K3
        ret, _ = e.returns()
        d_ret = torch.from_numpy(ret).to(dev)
        out.fill_(-1.0)
        eng.argmin_noalloc(d_ret.data_ptr(), N, out.data_ptr(), sample_offset=offset, device=0, stream=stream)
        got = out.cpu().numpy()
        assert got[0] == wr, (call, got, wr)                            # bit-exact, as every argmin test here
        assert int(got[1]) == offset + w, (call, got, w)
        assert w == int(np.argmin(ret))
    # ties go to the lowest index, NaN never wins, an all-NaN batch picks the first sample (K3's rules)
    t = torch.full((3000,), 5.0, dtype=torch.float64, device=dev); t[17] = float("nan"); t[1234] = 1.0; t[2999] = 1.0
    eng.argmin_noalloc(t.data_ptr(), 3000, out.data_ptr(), device=0, stream=stream)
    assert out.cpu().tolist() == [1.0, 1234.0]
    t.fill_(float("nan"))
    eng.argmin_noalloc(t.data_ptr(), 3000, out.data_ptr(), sample_offset=40, device=0, stream=stream)
    assert out.cpu().tolist()[1] == 40.0
    e.close()


@pytest.mark.gpu
def test_noalloc_rejects_oversize():
    import torch
    from mujoco_mpc_b200 import engine as eng
    dev = torch.device("cuda", 0)
    t = torch.zeros(8, dtype=torch.float64, device=dev); out = torch.zeros(2, dtype=torch.float64, device=dev)
    with pytest.raises(eng.EngineError, match="exceeds PS_NOALLOC_MAX_RETURNS"):
        eng.argmin_noalloc(t.data_ptr(), (1 << 20) + 1, out.data_ptr())


def test_noalloc_scratch_is_modest():
    from mujoco_mpc_b200 import build, engine as eng
    build.build()
    assert 0 < eng.lib().ps_noalloc_scratch_bytes() <= 64 * 1024
