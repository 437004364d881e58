"""Generates tests/golden/*: run from the repo root, `python tests/golden/make_golden.py`.

The reference cannot be built or imported here (needs CGAL 4.2 / Eigen / Boost), so the
golden vectors are (a) the numbers the reference publishes in its README and (b) values produced by the
CPU oracle cross-checked against exact rational arithmetic (oracle/pyoracle.py) at generation time.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from oracle import pyoracle  # noqa: E402
from oracle.oracle import OracleMesh  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

known = {
    "source": "/root/reference/README.md:126-131 (transcript of src/delaunay_2_test2.cpp-style program)",
    "regular_grid_sse": {"value": "5.800002", "line": "README.md:130", "status": "reproduced by the oracle"},
    "adaptive_grid_sse": {"value": "0.044019", "line": "README.md:131",
                          "status": "reproduced by the native host triangulation + adaptive queue (dli_di2) with the "
                                    "oracle's interp (tests/test_hostmesh.py); SURVEY.md F7's restatement did not"},
    "test2_variant_regular_grid_sse": {"value": "1.909975", "status": "unpublished; predicted by the restatement"},
}
with open(os.path.join(HERE, "readme_known_answers.json"), "w") as fh:
    json.dump(known, fh, indent=1)

c, f, t = helpers.grid_mesh(0.0)
q = helpers.grid_queries(0.0)
om = OracleMesh(c, f, t)
ys, s, st = om.interp(q, mode=0)
csv, _ = om.simplices()
ys_exact = np.array([float(pyoracle.interp_exact(c.tolist(), f.tolist(), csv.tolist(), p.tolist())[0]) for p in q])
assert np.max(np.abs(ys - ys_exact)) < 1e-13
np.savez_compressed(os.path.join(HERE, "config1_grid.npz"), coords=c, f=f, simplices=t, queries=q, ys=ys,
                    ys_exact=ys_exact, simplex=s)
print("sse", "%.6f" % float(np.sum((helpers.cobb_douglas(q) - ys) ** 2)))
