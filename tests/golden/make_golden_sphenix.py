#!/usr/bin/env python
"""Generates the SPHENIX golden fixtures tests/golden/sphenix/*.npz by running the REFERENCE's own
compiled operators of the three-loop scheme (oracle/_ref/libswiftref_sphenix.so: SPHENIX_SPH +
EOS_IDEAL_GAS, built from /root/reference by oracle/Makefile.ref VARIANT=sphenix):
sort -> density -> ghost -> GRADIENT loop -> extra ghost -> force -> end_force.

Run in the build container (where /root/reference exists):
    make -C oracle ref && python tests/golden/make_golden_sphenix.py

Each file holds the inputs (cell order, with the per-particle viscosity / diffusion state the scheme
carries from step to step), the tree and task list, and for every loop (0 first density pass, 1 ghost
redo sweeps, 2 force, 3 gradient) the interaction count, the per-particle neighbour counts and a SHA-256
of the sorted ordered pair list, plus every per-particle field after each stage.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)

import oracle_lib as ol  # noqa: E402
from make_golden import cells_to_arrays, ngb_counts, pair_digest  # noqa: E402
from swift_subtask_dev_b200 import ics  # noqa: E402

OUT_DIR = os.path.join(HERE, "sphenix")

IN_STATE = (("in_rho", "g_rho"), ("in_pressure", "g_pressure"), ("in_f", "g_f"), ("in_soundspeed", "g_soundspeed"),
            ("in_balsara", "g_balsara"), ("in_alpha_visc", "e_alpha_visc"), ("in_alpha_diff", "e_alpha_diff"),
            ("in_div_v_prev", "e_div_v_prev"))


def rough(ic, seed):
    """Gives the scheme something to do: a spread of internal energies (thermal diffusion, laplace_u), a
    converging velocity field on top of the noise (div_v < 0: the viscosity source term) and a
    viscosity / diffusion state as a previous step would have left it."""
    rng = np.random.default_rng(seed)
    n = len(ic["h"])
    ic["u"] = (ic["u"] * (1.0 + 0.5 * rng.random(n))).astype(np.float32)
    centre = ic["x"].mean(axis=0)
    ic["v"] = (ic["v"] - 0.3 * (ic["x"] - centre)).astype(np.float32)
    ic["in_alpha_visc"] = (0.1 + 0.5 * rng.random(n)).astype(np.float32)
    ic["in_alpha_diff"] = (0.3 * rng.random(n)).astype(np.float32)
    ic["in_div_v_prev"] = rng.standard_normal(n).astype(np.float32)
    return ic


def warm_start(ic, kw):
    """One all-active step with the reference; returns ic (cell order) with h and the carried state of it."""
    prm = ol.make_params(scheme=2, **kw)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    icc = ics.permute(ic, order)
    r = ol.run_step("ref", prm, icc, cells, tasks, nt, sphenix=ol.SPHENIX_PROPS)
    icc["h"] = r["g_h"].copy()
    for k_in, k_out in IN_STATE:
        icc[k_in] = r[k_out].copy()
    n = len(icc["h"])
    rng = np.random.default_rng(77)
    icc["x"] = np.mod(icc["x"] + rng.uniform(-0.1, 0.1, (n, 3)) / n ** (1.0 / 3.0), 1.0)
    return icc


def cases():
    c = {}
    kw = dict(periodic=True, alpha=0.1)
    c["uniform14_periodic"] = (rough(ics.uniform_random(14 ** 3, seed=3), 5), kw)
    c["clustered3k_open"] = (rough(ics.clustered(3000, nclumps=3, seed=9), 6), dict(periodic=False, alpha=0.1))
    ic = warm_start(rough(ics.uniform_random(11 ** 3, seed=21), 7), kw)
    ic["time_bin"] = np.random.default_rng(1).integers(1, 4, len(ic["h"])).astype(np.int8)
    c["random11_timebins"] = (ic, dict(periodic=True, alpha=0.1, max_active_bin=2))
    # strays that reach h_max without a neighbour: hydro_part_has_no_neighbours, then the gradient / force loops
    # with nothing to sum for them
    ic = clump = ics.clump_and_strays()
    n = len(ic["h"])
    rng = np.random.default_rng(8)
    ic["in_alpha_visc"] = (0.1 + 0.5 * rng.random(n)).astype(np.float32)
    ic["in_alpha_diff"] = (0.3 * rng.random(n)).astype(np.float32)
    ic["in_div_v_prev"] = rng.standard_normal(n).astype(np.float32)
    c["strays640_open"] = (ic, dict(periodic=False, alpha=0.1, h_max=0.045))
    return c


INTEG_KEYS = ("alpha_visc", "alpha_diff", "div_v_prev", "div_v_dt", "laplace_u")


def integ_case():
    """Engine steps with individual time bins in the three-loop scheme (tests/integ_lib.py: run_loop): drift, the
    seven hydro phases on the active particles, kick2, time-step, kick1; the alphas evolve over dt_alpha of each
    particle's own bin."""
    import integ_lib as il
    kw = dict(periodic=True, scheme=2, alpha=0.1)
    ip = il.default_iparams(scheme=2)
    ic = ics.clustered(4000, nclumps=4, seed=12)
    rng = np.random.default_rng(4)
    ic["u"] = (ic["u"] * (1.0 + 0.5 * rng.random(len(ic["h"])))).astype(np.float32)
    icc, tree = il.prepare(ic, kw)
    return icc, tree, kw, ip, 6


def main_integ():
    import integ_lib as il
    icc, tree, kw, ip, nsteps = integ_case()
    s, log = il.run_loop("ref", icc, kw, ip, tree, nsteps, variant="sphenix", sphenix=ol.SPHENIX_PROPS)
    so, lo = il.run_loop("oracle", icc, kw, ip, tree, nsteps, sphenix=ol.SPHENIX_PROPS)
    ties = [k for k in il.STATE_KEYS + INTEG_KEYS if not np.array_equal(so[k], s[k])]
    if ties or lo != log:
        raise SystemExit("integ: oracle and reference differ in %s (sort-key tie?): pick another seed" % ties)
    z = {k: s[k] for k in il.STATE_KEYS + INTEG_KEYS}
    z["log"] = np.array(log, dtype=np.float64)
    z["log_int"] = np.array([[l[0], l[1], l[2], l[5]] for l in log], dtype=np.int64)
    path = os.path.join(OUT_DIR, "integ", "clustered4k_bins.npz")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    np.savez_compressed(path, **z)
    print("integ clustered4k_bins: steps", nsteps, "bins", np.bincount(s["time_bin"])[40:], "updated", [l[2] for l in log],
          "alpha_visc range", float(s["alpha_visc"].min()), float(s["alpha_visc"].max()))


def main():
    if not ol.ref_available("sphenix"):
        sys.exit("oracle/_ref/libswiftref_sphenix.so is not built")
    os.makedirs(OUT_DIR, exist_ok=True)
    for name, (ic, kw) in cases().items():
        prm = ol.make_params(record=True, scheme=2, **kw)
        order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
        icc = ics.permute(ic, order)
        r = ol.run_step("ref", prm, icc, cells, tasks, nt, record=True, sphenix=ol.SPHENIX_PROPS)
        loc, width, ints = cells_to_arrays(cells)
        out = dict(x=icc["x"], v=icc["v"], mass=icc["mass"], h=icc["h"], u=icc["u"], id=icc["id"],
                   cell_loc=loc, cell_width=width, cell_ints=ints, cdim=np.array(cdim, dtype=np.int32),
                   tasks=np.array([[tasks[k].type, tasks[k].ci, tasks[k].cj] for k in range(nt)],
                                  dtype=np.int32).reshape(nt, 3),
                   counts=np.array(r["counts"], dtype=np.int64),
                   digests=np.array([pair_digest(r["pairs"][s]) for s in range(4)]),
                   params=np.array(repr(kw)), props=np.array(repr(ol.SPHENIX_PROPS)), variant=np.array("sphenix"))
        if icc.get("time_bin") is not None:
            out["time_bin"] = icc["time_bin"]
        for k_in, _ in IN_STATE:
            if icc.get(k_in) is not None:
                out[k_in] = icc[k_in]
        for s in range(4):
            out["ngb%d" % s] = ngb_counts(r["pairs"][s], icc["id"])
        for k in list(ol.OUT_FIELDS) + list(ol.SPHENIX_OUT) + ["f_min_ngb_time_bin"]:
            out[k] = r[k]
        path = os.path.join(OUT_DIR, name + ".npz")
        np.savez_compressed(path, **out)
        print("%-22s n=%d cells=%d tasks=%d counts=%s -> %d bytes" %
              (name, len(icc["h"]), len(cells), nt, r["counts"], os.path.getsize(path)))


if __name__ == "__main__":
    main()
    main_integ()
