"""Writes tests/golden/integ/*.npz: the state the REFERENCE's own compiled cell_drift_part, runner_do_kick1/2,
runner_do_timestep and hydro operators (oracle/_ref, built from /root/reference by oracle/Makefile.ref) leave
after a few engine steps with individual time bins (tests/integ_lib.py: run_loop).  Run in the build
container (needs /root/reference); the fixtures travel to the GPU box.

    python tests/golden/make_golden_integ.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import integ_lib as il  # noqa: E402
import oracle_lib as ol  # noqa: E402
from swift_subtask_dev_b200 import ics  # noqa: E402

CASES = {
    # name: (ic factory, step params, integrator params, scheme variant, steps)
    "clustered4k_bins": (lambda: ics.clustered(4000, nclumps=4, seed=11), dict(periodic=True),
                         il.default_iparams(), "minimal", 6),
    "sedov12_big_steps": (lambda: ics.lattice(12, jitter=0.05, sedov=True), dict(periodic=True),
                          dict(il.default_iparams(cfl=1.5, dt_max=0.1), log_max_h_change=8.0), "minimal", 4),
    "impact3k_planetary": (lambda: ics.two_body_impact(3000, seed=5), None, None, "planetary", 4),
}


def case(name):
    make, kw, ip, variant, nsteps = CASES[name]
    ic = make()
    eos = None
    if variant == "planetary":
        kw = dict(periodic=False, dim=tuple(float(v) for v in ic["dim"]), scheme=1, alpha=1.5, h_max=float(ic["h_max"]))
        ip = il.default_iparams(scheme=1, h_max=float(ic["h_max"]), dt_max=1e3, time_end=1e5)
        eos = ol.eos_from_json()
    icc, tree = il.prepare(ic, kw)
    return icc, tree, kw, ip, variant, nsteps, eos


def main():
    out_dir = os.path.join(HERE, "integ")
    os.makedirs(out_dir, exist_ok=True)
    for name in CASES:
        icc, tree, kw, ip, variant, nsteps, eos = case(name)
        s, log = il.run_loop("ref", icc, kw, ip, tree, nsteps, eos=eos, variant=variant)
        # a fixture must not hinge on the order of two particles with the SAME sort key (unspecified in the
        # reference, runner_sort.c is not stable): the restatement has to land on the same bits
        so, _ = il.run_loop("oracle", icc, kw, ip, tree, nsteps, eos=eos, variant=variant)
        ties = [k for k in il.STATE_KEYS if not np.array_equal(so[k], s[k])]
        if ties:
            raise SystemExit("%s: oracle and reference differ in %s (sort-key tie?): pick another seed" % (name, ties))
        z = {k: s[k] for k in il.STATE_KEYS}
        z["log"] = np.array(log, dtype=np.float64)
        z["log_int"] = np.array([[l[0], l[1], l[2], l[5]] for l in log], dtype=np.int64)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **z)
        print(name, "steps", nsteps, "bins", np.bincount(s["time_bin"])[40:], "updated", [l[2] for l in log])


if __name__ == "__main__":
    main()
