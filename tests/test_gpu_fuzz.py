"""GPU: a short run of the randomised differential driver (scripts/fuzz_parity.py): random meshes in 2D / 3D / 4D, random
engine options (walk-record layouts, start copies, rings, epilogues), coordinate scales and shifts, sign-changing and
vanishing data, adversarial queries -- against the CPU oracle, same bars as the parity tests."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("seed", [11, 12])
def test_randomised_differential_run(gpu, seed):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "fuzz_parity.py"), "25", str(seed)],
                       capture_output=True, text=True, timeout=600)
    last = json.loads(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0 and last["failures"] == 0 and last["iterations"] >= 10, r.stdout[-3000:] + r.stderr[-2000:]
    assert last["worst_rel"] <= 1e-12
