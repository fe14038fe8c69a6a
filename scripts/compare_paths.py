"""Runs the same step through the first-generation kernels (SWIFT_CUDA_PATH=0) and through a set
of leaf-resident kernels (path bits: 1 density, 2 fused ghost, 4 force) and compares every
output: counters must be equal; fields bit-equal except what the FAST force arithmetic touches
(a, u_dt, h_dt, v_sig), which are reported as relative errors against the scale of the field.

  python scripts/compare_paths.py --side 64 --path 7 [--workload clustered|uniform|lattice]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def run(ic, path, prm_kw, reps=1, fields=0):
    os.environ["SWIFT_CUDA_PATH"] = str(path)
    os.environ["SWIFT_CUDA_FIELDS"] = str(fields)
    from swift_subtask_dev_b200 import HydroBackend, make_params
    be = HydroBackend(make_params(**prm_kw))
    be.build(ic)
    best = None
    for r in range(reps):
        be.step(do_sort=True, restore=True)
        t = be.timings()
        if best is None or t["total_ms"] < best["total_ms"]:
            best = t
    out = be.results(first_pass=True)
    cnt = be.counters()
    be.close()
    return out, cnt, best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--side", type=int, default=64)
    ap.add_argument("--path", type=int, default=1)
    ap.add_argument("--workload", default="uniform")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--open", action="store_true")
    ap.add_argument("--fast", action="store_true")
    a = ap.parse_args()
    from swift_subtask_dev_b200 import ics
    n = a.side ** 3
    if a.workload == "uniform":
        ic = ics.uniform_random(n, seed=42)
    elif a.workload == "clustered":
        ic = ics.clustered(n, seed=7)
    else:
        ic = ics.lattice(a.side, jitter=0.1, seed=1234)
    kw = dict(periodic=not a.open)
    o0, c0, t0 = run(ic, 0, kw, a.reps)
    o1, c1, t1 = run(ic, a.path, kw, a.reps, 1 if a.fast else 0)
    res = {"n": n, "path": a.path, "t_old": t0, "t_new": t1}
    ok = True
    for k in ("n_density", "n_density_redo", "n_force", "ghost_sweeps"):
        res[k] = (c0[k], c1[k])
        ok &= c0[k] == c1[k]
    soft = {"a", "u_dt", "h_dt", "v_sig"} if ((a.path & 4) or a.fast) else set()
    for k in o0:
        x, y = np.asarray(o0[k]), np.asarray(o1[k])
        nbad = int(np.count_nonzero((x != y) & ~(np.isnan(x.astype(np.float64)) & np.isnan(y.astype(np.float64)))))
        if k in soft:
            scale = float(np.sqrt(np.mean(x.astype(np.float64) ** 2))) or 1.0
            err = float(np.max(np.abs(x.astype(np.float64) - y.astype(np.float64)))) / scale
            rel = np.abs(x.astype(np.float64) - y.astype(np.float64)) / np.maximum(np.abs(x.astype(np.float64)), 1e-3 * scale)
            res[k] = {"differ": nbad, "max_abs_over_rms": err, "max_rel_floor1e-3rms": float(rel.max())}
        else:
            res[k] = nbad
            ok &= nbad == 0
    res["exact_ok"] = bool(ok)
    print(json.dumps(res))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
