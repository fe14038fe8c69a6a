"""Diagnostic (GPU box): every field of every stage of the SPHENIX path against the golden fixtures, all mismatches
listed instead of stopping at the first.  python scripts/sphenix_diag.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as gu  # noqa: E402
import test_zy_gpu_sphenix as T  # noqa: E402


def report(tag, got, ref, table, act):
    for kg, ko in table:
        a, b = np.asarray(got[kg]), np.asarray(ref[ko])
        if act is not None:
            a, b = a[act], b[act]
        bad = (a != b) & ~(np.isnan(a.astype(np.float64)) & np.isnan(b.astype(np.float64)))
        nb = int(np.count_nonzero(bad.reshape(len(a), -1).any(axis=1)))
        if nb:
            d = np.abs(a.astype(np.float64) - b.astype(np.float64))
            rel = d / np.maximum(np.abs(b.astype(np.float64)), 1e-30)
            i = int(np.argmax(d.reshape(len(a), -1).max(axis=1)))
            print("  %-14s %-20s MISMATCH %6d / %d  max abs %.3e  max rel %.3e  e.g. [%d] got %s ref %s" %
                  (tag, kg, nb, len(a), d.max(), rel.max(), i, a[i], b[i]))
        else:
            print("  %-14s %-20s ok" % (tag, kg))


for name in gu.sphenix_names():
    print("==", name)
    g = gu.load_sphenix(name)
    act = T._active(g)
    be = T._build_golden(g, record=True)
    be.runner_do_hydro_sort()
    be.runner_do_density()
    be.runner_do_ghost()
    report("ghost", T._both(be), g, T.AFTER_GHOST, act)
    be.runner_do_gradient()
    report("gradient", T._both(be), g, T.AFTER_GRADIENT, act)
    be.runner_do_extra_ghost()
    report("extra_ghost", T._both(be), g, T.AFTER_EXTRA_GHOST, act)
    be.runner_do_force()
    out = T._both(be)
    report("force", out, g, T.AFTER_FORCE, act)
    c = be.counters()
    print("  counts", [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]], "ref",
          [int(v) for v in g["counts"]])
    for s in range(4):
        p = be.fetch_pairs(s, len(g["h"]) * 400)
        print("  pairs loop %d: %s" % (s, "ok" if gu.pair_digest(p) == str(g["digests"][s]) else "DIFFER (%d)" % len(p)))
    be.close()
