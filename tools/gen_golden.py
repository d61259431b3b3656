#!/usr/bin/env python3
"""Generate tests/golden/norm_spline_reference.json from the reference's OWN norm.cc / spline.cc
(compiled unmodified into oracle/_ref by oracle/Makefile).  Run in the build container, where
/root/reference exists; the committed JSON travels to the GPU box, the reference does not."""
import json
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402


def main():
    pyoracle.build()
    assert pyoracle.ref_lib() is not None, "oracle/_ref not built (reference tree absent?)"
    rng = np.random.default_rng(20260411)
    out = {"source": "mjpc/norm.cc + mjpc/spline/spline.cc of the reference, compiled unmodified (oracle/Makefile: ref)",
           "norm": [], "spline": []}
    # norm_test.cc:120-134 parameterisation + extra points
    norm_cases = [(0, [0.1]), (1, [0.1, 2.0]), (1, [0.3, 1.5]), (2, [0.1]), (3, [0.1]), (5, [2.0]), (5, [1.5]), (6, [0.1]),
                  (7, [0.1, 2.0]), (7, [0.2, 3.0]), (8, [0.1]), (-1, [])]
    points = [[0, 0], [1, 0], [-1, 0], [1, 1], [-1, -1]]             # norm_test.cc:38-40
    for ntype, params in norm_cases:
        pts = [p for p in points] + [list(rng.normal(0, 0.7, n)) for n in (1, 2, 3, 5, 12)]
        for x in pts:
            if ntype == -1:
                x = x[:1]
            out["norm"].append({"type": ntype, "params": params, "x": list(map(float, x)),
                                "value": pyoracle.ref_norm(x, params, ntype)})
    for interp in (0, 1, 2):
        for P, dim in ((1, 1), (2, 2), (3, 1), (5, 2), (10, 1), (3, 12)):
            times = np.cumsum(rng.uniform(0.05, 0.4, P)) + rng.uniform(-1, 1)
            knots = rng.normal(0, 1, (P, dim))
            ts = list(np.linspace(times[0] - 0.3, times[-1] + 0.3, 23)) + list(times)     # incl. exact node times
            for t in ts:
                out["spline"].append({"interp": interp, "times": list(map(float, times)), "knots": knots.tolist(), "t": float(t),
                                      "value": pyoracle.ref_spline_sample(times, knots, interp, t).tolist()})
    path = os.path.join(ROOT, "tests", "golden", "norm_spline_reference.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, len(out["norm"]), "norm and", len(out["spline"]), "spline vectors")


if __name__ == "__main__":
    main()
