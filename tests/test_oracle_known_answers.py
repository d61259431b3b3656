"""Pins the oracle's cost side against the reference: (1) golden vectors produced by the
reference's own norm.cc / spline.cc compiled unmodified (tests/golden/norm_spline_reference.json,
generator tools/gen_golden.py); (2) the known answers of the reference's gtest suites
(mjpc/test/spline/spline_test.cc:115-160, mjpc/test/tasks/task_test.cc:49-98); (3) when
oracle/_ref is present (build container), the live reference library directly."""
import json
import math
import os
import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(HERE, "golden", "norm_spline_reference.json")) as f:
        return json.load(f)


def test_norm_matches_reference_golden(oracle_lib, golden):
    for g in golden["norm"]:
        v = oracle_lib.norm(g["x"], g["params"], g["type"])
        assert v == pytest.approx(g["value"], rel=1e-14, abs=1e-15), g


def test_spline_matches_reference_golden(oracle_lib, golden):
    for g in golden["spline"]:
        v = oracle_lib.spline_sample(g["times"], g["knots"], g["interp"], g["t"])
        assert np.allclose(v, g["value"], rtol=1e-14, atol=1e-15), g


def test_live_reference_library_when_present(oracle_lib):
    if oracle_lib.ref_lib() is None:
        pytest.skip("oracle/_ref not present on this machine (reference tree absent)")
    rng = np.random.default_rng(3)
    for _ in range(200):
        P, dim, interp = int(rng.integers(1, 8)), int(rng.integers(1, 4)), int(rng.integers(0, 3))
        times = np.cumsum(rng.uniform(0.01, 1, P)); knots = rng.normal(0, 1, (P, dim)); t = rng.uniform(-0.5, times[-1] + 0.5)
        assert np.allclose(oracle_lib.spline_sample(times, knots, interp, t), oracle_lib.ref_spline_sample(times, knots, interp, t),
                           rtol=1e-14, atol=1e-15)
    for ntype, params in [(0, []), (1, [0.2, 1.7]), (2, [0.04]), (3, [0.3]), (5, [1.3]), (6, [0.01]), (7, [0.1, 2.5]), (8, [0.2])]:
        for _ in range(50):
            x = rng.normal(0, 1, int(rng.integers(1, 9)))
            assert oracle_lib.norm(x, params, ntype) == pytest.approx(oracle_lib.ref_norm(x, params, ntype), rel=1e-14, abs=1e-15)


# ---- spline_test.cc known answers
def test_spline_zero_linear_cubic_known_answers(oracle_lib):
    t, k = [1.0, 2.0], [[1.0, 2.0], [3.0, 4.0]]
    assert list(oracle_lib.spline_sample(t, k, 0, 1.5)) == [1.0, 2.0]          # ZeroOrder, :115-122
    assert list(oracle_lib.spline_sample(t, k, 1, 1.5)) == [2.0, 3.0]          # Linear, :124-131
    assert list(oracle_lib.spline_sample(t, k, 2, 1.5)) == [2.0, 3.0]          # Cubic, :133-139
    t4, k4 = [0.0, 1.0, 2.0, 3.0], [[1.0, 2.0], [1.0, 2.0], [3.0, 4.0], [3.0, 4.0]]
    assert np.allclose(oracle_lib.spline_sample(t4, k4, 2, 1.5), [2.0, 3.0], atol=1e-15)   # :141-147
    for x in np.arange(0.0, 1.0 + 1e-9, 0.125):                                # :149-159, y = -x^3 + 2x^2
        y = oracle_lib.spline_sample([-1.0, 0.0, 1.0], [[1.0], [0.0], [1.0]], 2, x)[0]
        assert y == pytest.approx(-x ** 3 + 2 * x ** 2, abs=1e-15)


def test_spline_holds_outside_range_and_empty(oracle_lib):
    t, k = [1.0, 2.0, 3.0], [[1.0], [2.0], [3.0]]
    for interp in (0, 1, 2):
        assert oracle_lib.spline_sample(t, k, interp, 0.0)[0] == 1.0           # DiscardBefore test: Sample(0.0) holds first
        assert oracle_lib.spline_sample(t, k, interp, 9.0)[0] == 3.0
    assert oracle_lib.spline_sample(np.zeros(0), np.zeros((0, 1)), 2, 0.5)[0] == 0.0  # empty spline samples 0 (spline.cc:108-111)


# ---- task_test.cc known answers (particle_task.xml cost spec restated as a PsTask)
def test_cost_terms_and_risk_known_answers(oracle_lib):
    from mujoco_mpc_b200 import mjcf, task
    xml = """<mujoco><custom><numeric name="task_risk" data="1.0"/><numeric name="residual_dummy1" data="0.05"/>
      <numeric name="residual_dummy2" data="-0.1"/></custom>
      <worldbody><body name="b"><joint type="slide" axis="1 0 0"/><geom size=".01" mass="1"/></body></worldbody>
      <sensor><user name="Position" dim="2" user="0 5.0 0.0 10.0"/><user name="Velocity" dim="2" user="0 0.1 0.0 1.0"/></sensor></mujoco>"""
    m = mjcf.compile_xml(xml)
    t = task.make_task(m, "ParticleTest")
    assert t.risk == pytest.approx(1.0) and t.num_parameters == 2
    assert t.parameters[0] == pytest.approx(0.05) and t.parameters[1] == pytest.approx(-0.1)
    assert (t.num_residual, t.num_term) == (4, 2)
    assert list(t.dim_norm_residual[:2]) == [2, 2] and list(t.num_norm_parameter[:2]) == [0, 0] and list(t.norm[:2]) == [0, 0]
    assert t.weight[0] == pytest.approx(5.0) and t.weight[1] == pytest.approx(0.1)
    o = oracle_lib.Oracle(m, t)
    r = np.array([1.0e-3, 2.0e-3, 3.0e-3, 4.0e-3])
    c = 5.0 * 0.5 * r[:2] @ r[:2] + 0.1 * 0.5 * r[2:] @ r[2:]
    assert o.cost_terms(r).sum() == pytest.approx(c, abs=1e-15)
    t.risk = 0.2
    o.set_task(t)
    assert o.cost(r) == pytest.approx((math.exp(0.2 * c) - 1.0) / 0.2, abs=1e-15)
    t.risk = 0.0
    o.set_task(t)
    assert o.cost(r) == pytest.approx(c, abs=1e-15)


def test_task_parsing_rejects_bad_specs():
    from mujoco_mpc_b200 import mjcf, task
    xml = """<mujoco><worldbody><body><joint type="slide"/><geom size=".01" mass="1"/></body></worldbody>
      <sensor><user name="A" dim="1" user="6 1.0 0 1"/></sensor></mujoco>"""     # smooth-abs needs a parameter
    with pytest.raises(ValueError):
        task.make_task(mjcf.compile_xml(xml), "ParticleTest")
