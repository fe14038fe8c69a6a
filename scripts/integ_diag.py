"""Diagnostic: per step, which fields of the device-resident loop differ from the oracle's (tests/integ_lib.py)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import integ_lib as il
import test_integrate as ti

name = sys.argv[1] if len(sys.argv) > 1 else "impact3k_planetary"
icc, tree, kw, ip, variant, nsteps, eos = ti._case(name)
ref = []
so, lo = il.run_loop("oracle", icc, kw, ip, tree, nsteps, eos=eos, variant=variant,
                     on_step=lambda step, s: ref.append({k: s[k].copy() for k in il.STATE_KEYS}))
be = ti._gpu_backend(kw, eos)


def check(step, g):
    for k in il.STATE_KEYS:
        a, b = np.asarray(g[k]).reshape(len(g["h"]), -1), ref[step][k].reshape(len(g["h"]), -1)
        bad = np.nonzero((a != b).any(axis=1))[0]
        if len(bad):
            i = bad[0]
            print("step", step, k, "differs on", len(bad), "particles; first", i, a[i], b[i], "bin", ref[step]["time_bin"][i],
                  "mat", None if icc.get("mat_id") is None else icc["mat_id"][i])


sg, lg = il.run_loop_gpu(be, icc, ip, tree, nsteps, on_step=check)
print("gpu log", [l[:3] for l in lg])
print("ora log", [l[:3] for l in lo])
