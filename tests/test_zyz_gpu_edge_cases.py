"""Degenerate inputs on the GPU (no particle, one, two, nobody active, one in a hundred active) through the library's
own entry points, against the oracle: scripts/edge_cases_gpu_check.py in a process of its own.

STATUS: written after the round's GPU budget had been spent; the entry points are the ones every other GPU test
uses, the inputs are new.  xfail(strict=False) until it has been seen on hardware once: XPASS = all cases equal to
the oracle, XFAIL = the script's output names the case."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.xfail(strict=False, reason="first run on hardware (written after the round's GPU budget was spent)")
def test_degenerate_inputs_against_the_oracle():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "edge_cases_gpu_check.py")], cwd=ROOT,
                       capture_output=True, text=True, timeout=300)
    sys.stdout.write(r.stdout[-6000:])
    sys.stderr.write(r.stderr[-3000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "all ok" in r.stdout
