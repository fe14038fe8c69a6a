"""Writes tests/golden/limiter/*.npz: the time-step limiter (swift --limiter) as the REFERENCE's own compiled
operators do it (oracle/_ref, built from /root/reference by oracle/Makefile.ref):

  * loop_*:  the tasks of subtype limiter (runner_do{self,pair}1_branch_limiter, runner_dosub_*1_limiter,
             src/runner_doiact_limiter.c) on a particle set with given time bins: p->limiter_data.wakeup of
             every particle;
  * apply_*: runner_do_limiter / timestep_limit_part on a state in the middle of a run: the state afterwards;
  * run_*:   engine steps with individual time bins and the limiter switched on (tests/integ_lib.py: run_loop
             with limiter=True: ... kick2, timestep, limiter loop, runner_do_limiter, kick1): the final state
             and the per-step log (active, woken, woken inactive, re-kicked particles).

Run in the build container (needs /root/reference); the fixtures travel to the GPU box.

    python tests/golden/make_golden_limiter.py
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

OUT_DIR = os.path.join(HERE, "limiter")

# name: (ic factory, step params, max_active_bin, highest bin, seed of the bins)
LOOP_CASES = {
    # deep tree (split cells, sub tasks), open box, wide spread of h
    "loop_clustered3k_open": (lambda: ics.clustered(3000, nclumps=3, seed=9), dict(periodic=False), 3, 9, 4),
    # pairs through the periodic wrap, no split cells
    "loop_uniform12_periodic": (lambda: ics.uniform_random(12 ** 3, seed=5), dict(periodic=True), 2, 7, 6),
    # everybody starts a step (step 0 of a run): only the bins decide
    "loop_lattice10_all_starting": (lambda: ics.lattice(10, jitter=0.2, sedov=False), dict(periodic=True), 8, 8, 7),
}

# name: (ic factory, step params, integrator params, variant, steps)
RUN_CASES = {
    "run_sedov14": (lambda: ics.lattice(14, jitter=0.05, sedov=True), dict(periodic=True), il.default_iparams(),
                    "minimal", 12),
    "run_sedov12_sphenix": (lambda: ics.lattice(12, jitter=0.05, sedov=True), dict(periodic=True, scheme=2, alpha=0.1),
                            il.default_iparams(scheme=2), "sphenix", 10),
}


def loop_case(name):
    """Particles in cell order with converged h, the tree, and the bins 'after the time-step task': inactive
    particles (bin > max_active_bin) anywhere up to `top`, starting ones anywhere from 1 to max_active_bin."""
    make, kw, mab, top, seed = LOOP_CASES[name]
    icc, tree = il.prepare(make(), kw)
    n = len(icc["h"])
    rng = np.random.default_rng(seed)
    bins = rng.integers(1, top + 1, n).astype(np.int8)
    icc["time_bin"] = bins
    return icc, tree, dict(kw, max_active_bin=mab)


def run_case(name):
    make, kw, ip, variant, nsteps = RUN_CASES[name]
    icc, tree = il.prepare(make(), kw)
    return icc, tree, kw, ip, variant, nsteps


def run_loop_only(which, icc, tree, kw, variant="minimal"):
    cells, tasks, nt = tree
    prm = ol.make_params(**kw)
    ic = dict(icc, l_time_bin=icc["time_bin"], l_only=True)
    return ol.run_step(which, prm, ic, cells, tasks, nt, variant=variant)["l_wakeup"]


def apply_case(which="ref"):
    """A state in the middle of run_sedov14 (after its last kick1), the next sync point, and wake-up calls for
    a third of the particles, active and inactive ones: all three branches of timestep_limit_part in quantity."""
    icc, tree, kw, ip, variant, nsteps = run_case("run_sedov14")
    s, log = il.run_loop(which, icc, kw, ip, tree, nsteps, variant=variant, limiter=True)
    ti_current = log[-1][5]
    mab = il.get_max_active_bin(ti_current)
    n = len(s["h"])
    rng = np.random.default_rng(12)
    woken = rng.random(n) < 0.33
    new_bin = np.maximum(1, s["time_bin"].astype(np.int32) - rng.integers(1, 4, n))
    woken &= new_bin < s["time_bin"]
    wake = np.where(woken, 2 - new_bin, il.NOT_AWAKE).astype(np.int8)
    return s, ip, variant, ti_current, mab, wake


def main():
    os.makedirs(OUT_DIR, exist_ok=True)
    for name in LOOP_CASES:
        icc, tree, kw = loop_case(name)
        wr = run_loop_only("ref", icc, tree, kw)
        wo = run_loop_only("oracle", icc, tree, kw)
        assert np.array_equal(wr, wo), name + ": oracle and reference differ"
        np.savez_compressed(os.path.join(OUT_DIR, name + ".npz"), l_wakeup=wr, time_bin=icc["time_bin"], h=icc["h"],
                            x=icc["x"])
        woken = wr != il.NOT_AWAKE
        starting = icc["time_bin"] <= kw["max_active_bin"]
        print(name, "n", len(wr), "woken", int(woken.sum()), "of them starting", int((woken & starting).sum()),
              "wakeup values", sorted(set(wr[woken].tolist())))
    s, ip, variant, ti_current, mab, wake = apply_case("ref")
    census = il.limiter_census(wake, s["time_bin"], mab)
    before = {k: s[k].copy() for k in il.STATE_KEYS}
    chk = il.Checker("ref", ip, variant=variant)
    w = wake.copy()
    t_end, nlim = chk.limiter(s, w, ti_current, mab)
    assert (w == il.NOT_AWAKE).all()
    z = {"before_" + k: before[k] for k in il.STATE_KEYS}
    z.update({"after_" + k: s[k] for k in il.STATE_KEYS})
    z.update(wake=wake, scalars=np.array([ti_current, mab, t_end, nlim], dtype=np.int64),
             census=np.array(census, dtype=np.int64))
    np.savez_compressed(os.path.join(OUT_DIR, "apply_sedov14.npz"), **z)
    print("apply_sedov14: ti_current", ti_current, "max_active_bin", mab, "census (woken, inactive, re-kicked)", census,
          "ti_end_min", t_end)
    for name in RUN_CASES:
        icc, tree, kw, ip, variant, nsteps = run_case(name)
        sx = ol.SPHENIX_PROPS if variant == "sphenix" else None
        s, log = il.run_loop("ref", icc, kw, ip, tree, nsteps, variant=variant, sphenix=sx, limiter=True)
        so, lo = il.run_loop("oracle", icc, kw, ip, tree, nsteps, variant=variant, sphenix=sx, limiter=True)
        ties = [k for k in il.STATE_KEYS if not np.array_equal(so[k], s[k])]
        if ties or lo != log:
            raise SystemExit("%s: oracle and reference differ in %s (sort-key tie?): pick another seed" % (name, ties))
        z = {k: s[k] for k in il.STATE_KEYS}
        z["log"] = np.array([l[:6] for l in log], dtype=np.float64)
        z["log_int"] = np.array([[l[0], l[1], l[2], l[5], l[6], l[7], l[8]] for l in log], dtype=np.int64)
        np.savez_compressed(os.path.join(OUT_DIR, name + ".npz"), **z)
        print(name, "steps", nsteps, "updated", [l[2] for l in log], "woken / inactive / re-kicked",
              [l[6:] for l in log])


if __name__ == "__main__":
    main()
