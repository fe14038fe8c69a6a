"""Writes tests/golden/eos_tables/impact3k_tables.npz and impact3k_tables_run.npz: one whole step (sort, density,
ghost, force) and a few engine steps (drift, hydro, kick2, time-step, kick1) of a two-body impact whose materials
live in the synthetic tables of make_eos_tables.py -- SESAME-like iron (300), ANEOS-like forsterite (400), HM80-like
rock (202) and HHe (200), a custom table (900) -- next to Tillotson iron (100), by the REFERENCE's own compiled
operators (oracle/_ref/libswiftref_planetary.so) with its own table loaders and interpolation.  Same format as
tests/golden/*.npz (make_golden.py) and tests/golden/integ/*.npz.  Run in the build container:

    python tests/golden/make_eos_tables.py && python tests/golden/make_golden_eos_step.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import eos_lib as el  # noqa: E402
import integ_lib as il  # noqa: E402
import oracle_lib as ol  # noqa: E402
from make_eos_tables import TABLES  # noqa: E402
from make_golden import cells_to_arrays, ngb_counts, pair_digest  # noqa: E402
from swift_subtask_dev_b200 import ics  # noqa: E402

OUT_DIR = os.path.join(HERE, "eos_tables")


def case():
    """Particles in cell order with converged h and the tree; the materials of ics.two_body_impact remapped."""
    ic = ics.two_body_impact(3000, seed=7)
    kw = dict(periodic=False, dim=tuple(float(v) for v in ic["dim"]), scheme=1, alpha=1.5, h_max=float(ic["h_max"]))
    ip = il.default_iparams(scheme=1, h_max=float(ic["h_max"]), dt_max=1e3, time_end=1e5)
    icc, tree = il.prepare(ic, kw)          # the smoothing lengths do not depend on the equation of state
    x = icc["x"]
    mat = np.asarray(icc["mat_id"]).copy()
    target = x[:, 0] < 0.5 * ic["dim"][0] + 1.0
    rng = np.random.default_rng(2)
    new = mat.copy()
    new[target & (mat == 100)] = 300                       # target core: SESAME-like iron
    new[target & (mat == 101)] = 400                       # target mantle: ANEOS-like forsterite
    new[~target & (mat == 101)] = 202                      # impactor mantle: HM80-like rock (bulk-modulus sound speed)
    pick = rng.random(len(mat))                            # impactor core stays Tillotson iron (100)
    new[target & (mat == 101) & (pick < 0.15)] = 200       # some HHe (ideal-gas-like sound speed from P)
    new[target & (mat == 101) & (pick > 0.9)] = 900        # some of the custom material
    icc["mat_id"] = new.astype(np.int32)
    return icc, tree, kw, ip


def load_reference_tables():
    for name, (fname, fmt, mat_id) in TABLES.items():
        el.ref_load(mat_id, os.path.join(OUT_DIR, fname))


def main():
    load_reference_tables()
    icc, tree, kw, ip = case()
    cells, tasks, nt = tree
    prm = ol.make_params(record=True, **kw)
    r = ol.run_step("ref", prm, icc, cells, tasks, nt, record=True, variant="planetary")
    loc, width, ints = cells_to_arrays(cells)
    out = dict(x=icc["x"], v=icc["v"], mass=icc["mass"], h=icc["h"], u=icc["u"], id=icc["id"], mat_id=icc["mat_id"],
               cell_loc=loc, cell_width=width, cell_ints=ints,
               tasks=np.array([[tasks[k].type, tasks[k].ci, tasks[k].cj] for k in range(nt)], dtype=np.int32).reshape(nt, 3),
               counts=np.array(r["counts"], dtype=np.int64), ghost_sweeps=np.int32(r["ghost_sweeps"]),
               digests=np.array([pair_digest(r["pairs"][s]) for s in range(3)]),
               params=np.array(repr(kw)), variant=np.array("planetary"))
    for s in range(3):
        out["ngb%d" % s] = ngb_counts(r["pairs"][s], icc["id"])
    for k in ol.OUT_FIELDS:
        out[k] = r[k]
    out["f_min_ngb_time_bin"] = r["f_min_ngb_time_bin"]
    np.savez_compressed(os.path.join(OUT_DIR, "impact3k_tables.npz"), **out)
    mats, cnt = np.unique(icc["mat_id"], return_counts=True)
    print("impact3k_tables: n", len(icc["h"]), "materials", dict(zip(mats.tolist(), cnt.tolist())), "counts", r["counts"],
          "P range %.3g..%.3g" % (r["g_pressure"].min(), r["g_pressure"].max()),
          "c range %.3g..%.3g" % (r["g_soundspeed"].min(), r["g_soundspeed"].max()))
    for m in mats:
        sel = icc["mat_id"] == m
        print("   mat", int(m), "P>0:", int((r["g_pressure"][sel] > 0).sum()), "of", int(sel.sum()),
              "median c %.3g" % np.median(r["g_soundspeed"][sel]))
    nsteps = 3
    s, log = il.run_loop("ref", icc, kw, ip, tree, nsteps, variant="planetary")
    z = {k: s[k] for k in il.STATE_KEYS}
    z["log"] = np.array(log, dtype=np.float64)
    z["log_int"] = np.array([[l[0], l[1], l[2], l[5]] for l in log], dtype=np.int64)
    np.savez_compressed(os.path.join(OUT_DIR, "impact3k_tables_run.npz"), **z)
    print("impact3k_tables_run: steps", nsteps, "bins", np.bincount(s["time_bin"])[30:], "updated", [l[2] for l in log])


if __name__ == "__main__":
    main()
