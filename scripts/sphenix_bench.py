"""One SPHENIX step (sort, density, ghost, gradient loop, extra ghost, force) on a uniform-random periodic box,
device-timed with the library's CUDA events, next to the two-loop Minimal scheme (EXACT fields mode) on the same
particles.  python scripts/sphenix_bench.py [side] -> one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from swift_subtask_dev_b200 import HydroBackend, ics, make_params  # noqa: E402
from swift_subtask_dev_b200 import backend as B  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 128
ic = ics.uniform_random(side ** 3, seed=42)
out = {"workload": "uniform-random periodic box %d^3" % side, "n": side ** 3}
schemes = (("minimal_exact", B.SCHEME_MINIMAL), ("sphenix", B.SCHEME_SPHENIX))
if len(sys.argv) > 2 and sys.argv[2] == "sphenix-only":  # under ncu
    schemes = schemes[1:]
nrep = 2 if len(schemes) == 1 else 4
for name, scheme in schemes:
    be = HydroBackend(make_params(periodic=True, alpha=0.1 if scheme == B.SCHEME_SPHENIX else 0.8, scheme=scheme))
    be.build(ic)
    best = None
    for rep in range(nrep):  # first repetition is the warm-up
        be.step(do_sort=True, restore=True)
        t = be.timings()
        if rep > 0 and (best is None or t["total_ms"] < best["total_ms"]):
            best = t
    c = be.counters()
    n_int = c["n_density"] + c["n_density_redo"] + c["n_force"]
    r = {"ms": best, "n_density": c["n_density"], "n_density_redo": c["n_density_redo"], "n_force": c["n_force"],
         "launches": c["n_kernel_launches"]}
    if scheme == B.SCHEME_SPHENIX:
        r["n_gradient"] = be.sphenix_results()["n_gradient"]
        n_int += r["n_gradient"]
        r["note"] = "ghost_ms includes the gradient loop and the extra ghost; repeated steps carry the evolved alphas"
    r["interactions_per_s"] = n_int / (best["total_ms"] * 1e-3)
    out[name] = r
    be.close()
print(json.dumps(out))
