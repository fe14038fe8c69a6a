#!/usr/bin/env python
"""Generates the golden fixtures tests/golden/*.npz by running the REFERENCE's
own compiled operators (oracle/_ref/libswiftref_*.so, built from
/root/reference by oracle/Makefile.ref) on small seeded inputs.

Run in the build container (where /root/reference exists):
    make -C oracle ref && python tests/golden/make_golden.py
The fixtures travel to the GPU box, where the reference does not exist.

Each file holds the inputs (cell order), the tree and task list, and for every
loop (0 first density pass, 1 ghost redo sweeps, 2 force) the interaction
count, the per-particle neighbour counts and a SHA-256 of the sorted ordered
pair list, plus every per-particle field after density / ghost / force.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as ol  # noqa: E402
from swift_subtask_dev_b200 import ics  # noqa: E402


def cases():
    c = {}
    c["lattice10_periodic"] = (ics.lattice(10, jitter=0.1, seed=1234), dict(periodic=True), "minimal")
    c["random12_open"] = (ics.uniform_random(12 ** 3, seed=9), dict(periodic=False), "minimal")
    c["clustered4k_periodic"] = (ics.clustered(4000, nclumps=4, seed=3), dict(periodic=True), "minimal")
    # mixed time bins: a SECOND step.  Step one (everything active) gives every particle
    # its h, rho, P, f, c, balsara; in step two only bins <= 2 are active and the inactive
    # particles carry the state of step one (in_* inputs).
    ic = ics.uniform_random(11 ** 3, seed=21)
    ic = warm_start(ic, dict(periodic=True), "minimal")
    ic["time_bin"] = np.random.default_rng(1).integers(1, 4, len(ic["h"])).astype(np.int8)
    c["random11_timebins"] = (ic, dict(periodic=True, max_active_bin=2), "minimal")
    ic = ics.two_body_impact(6000, seed=5)
    c["impact6k_planetary"] = (ic, dict(periodic=False, dim=ic["dim"], scheme=1, alpha=1.5, h_max=ic["h_max"]),
                               "planetary")
    return c


IN_STATE = (("in_rho", "g_rho"), ("in_pressure", "g_pressure"), ("in_f", "g_f"),
            ("in_soundspeed", "g_soundspeed"), ("in_balsara", "g_balsara"))


def warm_start(ic, kw, variant):
    """One all-active step with the reference; returns ic with h and the carried state of that step."""
    prm = ol.make_params(**kw)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    icc = ics.permute(ic, order)
    r = ol.run_step("ref", prm, icc, cells, tasks, nt, variant=variant)
    icc["h"] = r["g_h"].copy()
    for k_in, k_out in IN_STATE:
        icc[k_in] = r[k_out].copy()
    # a small drift, so that the h of the active particles has to be re-converged
    n = len(icc["h"])
    rng = np.random.default_rng(77)
    icc["x"] = np.mod(icc["x"] + rng.uniform(-0.1, 0.1, (n, 3)) / n ** (1.0 / 3.0), 1.0)
    return icc


def pair_digest(pairs):
    p = ol.sorted_pairs(np.asarray(pairs, dtype=np.int64))
    return hashlib.sha256(np.ascontiguousarray(p).tobytes()).hexdigest()


def ngb_counts(pairs, ids):
    """neighbour count of every particle, in the order of `ids`."""
    lut = {int(v): k for k, v in enumerate(ids)}
    out = np.zeros(len(ids), dtype=np.int32)
    if len(pairs):
        u, cnt = np.unique(pairs[:, 0], return_counts=True)
        for a, b in zip(u, cnt):
            out[lut[int(a)]] = b
    return out


def cells_to_arrays(cells):
    n = len(cells)
    loc = np.array([[c.loc[0], c.loc[1], c.loc[2]] for c in cells], dtype=np.float64).reshape(n, 3)
    width = np.array([[c.width[0], c.width[1], c.width[2]] for c in cells], dtype=np.float64).reshape(n, 3)
    ints = np.array([[c.first, c.count, c.parent, c.depth, c.split] + list(c.progeny) for c in cells],
                    dtype=np.int32).reshape(n, 13)
    return loc, width, ints


def main():
    for name, (ic, kw, variant) in cases().items():
        if not ol.ref_available(variant):
            print("skip", name, "(no compiled reference for", variant, ")")
            continue
        kw = dict(kw)
        ic.pop("dim", None)
        ic.pop("h_max", None)
        prm = ol.make_params(record=True, **kw)
        order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
        icc = ics.permute(ic, order)
        r = ol.run_step("ref", prm, icc, cells, tasks, nt, record=True, variant=variant)
        loc, width, ints = cells_to_arrays(cells)
        out = dict(x=icc["x"], v=icc["v"], mass=icc["mass"], h=icc["h"], u=icc["u"], id=icc["id"],
                   cell_loc=loc, cell_width=width, cell_ints=ints, cdim=np.array(cdim, dtype=np.int32),
                   tasks=np.array([[tasks[k].type, tasks[k].ci, tasks[k].cj] for k in range(nt)],
                                  dtype=np.int32).reshape(nt, 3),
                   counts=np.array(r["counts"], dtype=np.int64), ghost_sweeps=np.int32(r["ghost_sweeps"]),
                   digests=np.array([pair_digest(r["pairs"][s]) for s in range(3)]),
                   params=np.array(repr(kw)), variant=np.array(variant))
        if icc.get("mat_id") is not None:
            out["mat_id"] = icc["mat_id"]
        if icc.get("time_bin") is not None:
            out["time_bin"] = icc["time_bin"]
        for k_in, _ in IN_STATE:
            if icc.get(k_in) is not None:
                out[k_in] = icc[k_in]
        for s in range(3):
            out["ngb%d" % s] = ngb_counts(r["pairs"][s], icc["id"])
        for k in ol.OUT_FIELDS:
            out[k] = r[k]
        out["f_min_ngb_time_bin"] = r["f_min_ngb_time_bin"]
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print("%-24s n=%d cells=%d tasks=%d counts=%s sweeps=%d -> %d bytes" %
              (name, len(icc["h"]), len(cells), nt, r["counts"], r["ghost_sweeps"], os.path.getsize(path)))


if __name__ == "__main__":
    main()
