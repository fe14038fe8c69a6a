"""Degenerate inputs through the C ABI on the GPU, against the oracle where there is anything to compute:
python scripts/edge_cases_gpu_check.py

  no particle at all; one particle; two particles out of each other's reach; two within reach; a box in which
  nobody is active; a box of which one particle in a hundred is active (ragged units: most lanes idle)

The reference's own edge cases -- strays without neighbours (hydro_part_has_no_neighbours), cells with one particle,
every mix of active and inactive particles -- are tests/test_gpu_parity.py::test_strays_* and the reference's
testActivePair through the shim; this script holds the library's own entry points to the same kinds of input.
tests/test_zyz_gpu_edge_cases.py runs it in a process of its own."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import parity  # noqa: E402
from swift_subtask_dev_b200 import HydroBackend  # noqa: E402

FAILED = []


def check(ok, what):
    print(("ok   " if ok else "FAIL ") + what, flush=True)
    if not ok:
        FAILED.append(what)
    return ok


def particles(x, h=0.1, time_bin=None):
    x = np.asarray(x, dtype=np.float64).reshape(-1, 3)
    n = len(x)
    ic = {"x": x, "v": np.zeros((n, 3), np.float32), "mass": np.full(n, 1e-3, np.float32),
          "h": np.full(n, h, np.float32), "u": np.ones(n, np.float32), "id": np.arange(n, dtype=np.int64),
          "mat_id": None, "time_bin": None if time_bin is None else np.asarray(time_bin, dtype=np.int8)}
    return ic


def against_oracle(name, ic, kw):
    try:
        g = parity.run_gpu(ic, kw, record=True, pair_capacity=max(len(ic["h"]), 1) * 400)
    except Exception as e:  # noqa: BLE001
        check(False, "%s: the library raised %s" % (name, str(e)[:200]))
        return None
    o = parity.run_oracle_on_tree(ic, g["tree"], kw, record=True)
    act = None
    if ic.get("time_bin") is not None:
        act = ic["time_bin"] <= kw.get("max_active_bin", 56)
    r = parity.compare(g, o, active=act, verbose=False)
    try:
        parity.assert_exact(r)
        ok = all(parity.pair_set_diff(g["pairs"][k], o["pairs"][k]) == 0 for k in range(3))
        check(ok, "%s: counts %s, pair sets and every field of the active particles equal to the oracle's" % (
            name, [int(v) for v in o["counts"]]))
    except AssertionError as e:
        check(False, "%s: %s" % (name, str(e)[:300]))
    return g


def main():
    kw = dict(periodic=False, dim=(3.0, 3.0, 3.0), h_max=0.25)
    # nothing
    try:
        be = HydroBackend(parity.gpu_params(kw))
        be.build(particles(np.zeros((0, 3))))
        be.step()
        c = be.counters()
        be.close()
        check(c["n_density"] == 0 and c["n_force"] == 0, "no particle: build and step return, no interaction counted")
    except Exception as e:  # noqa: BLE001
        check(False, "no particle: the library raised %s" % str(e)[:200])
    # one, and two out of reach: h doubles up to h_max, nobody found (hydro_part_has_no_neighbours)
    g = against_oracle("one particle", particles([[1.5, 1.5, 1.5]]), kw)
    if g is not None:
        check(float(g["h"][0]) == np.float32(0.25), "one particle: h ends at h_max")
    against_oracle("two particles out of reach", particles([[0.5, 0.5, 0.5], [2.5, 2.5, 2.5]]), kw)
    against_oracle("two particles within reach", particles([[1.5, 1.5, 1.5], [1.55, 1.5, 1.5]]), kw)
    # nobody active / one in a hundred
    rng = np.random.default_rng(5)
    x = rng.random((4000, 3)) * 3.0
    tb = np.full(4000, 9, dtype=np.int8)
    against_oracle("nobody active", particles(x, h=0.25, time_bin=tb), dict(kw, h_max=0.4, max_active_bin=3))
    tb = np.where(rng.random(4000) < 0.01, 2, 9).astype(np.int8)
    against_oracle("one particle in a hundred active", particles(x, h=0.25, time_bin=tb),
                   dict(kw, h_max=0.4, max_active_bin=3))
    print("edge_cases_gpu_check:", "all ok" if not FAILED else "%d FAILED" % len(FAILED), flush=True)
    return 1 if FAILED else 0


if __name__ == "__main__":
    sys.exit(main())
