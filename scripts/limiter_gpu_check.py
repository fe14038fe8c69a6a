"""The time-step limiter on the GPU against the golden fixtures of tests/golden/limiter (the REFERENCE's own
compiled limiter loop and runner_do_limiter) and the C oracle:  python scripts/limiter_gpu_check.py [loop|run|all]

  loop  the limiter loop alone (swift_cuda_do_limiter_loop) on three particle sets with given time bins:
        p->limiter_data.wakeup of every particle, bit for bit; bins on the device and bins handed over
  run   engine steps with the limiter on, particles resident on the device (drift, hydro loops, kick2, time-step,
        limiter loop, runner_do_limiter, kick1): every field after EVERY step against the oracle, the final state
        and the per-step census (active, woken, woken inactive, re-kicked) against the reference's fixture

tests/test_zyz_gpu_limiter.py runs it in a process of its own (a device fault must not take the other GPU tests
with it).  Prints one line per check and exits 0 only when all pass."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import integ_lib as il  # noqa: E402
import oracle_lib as ol  # noqa: E402
import parity  # noqa: E402
import test_limiter_oracle as cpu  # noqa: E402
from swift_subtask_dev_b200 import HydroBackend  # noqa: E402
from swift_subtask_dev_b200 import backend as B  # noqa: E402

FAILED = []


def check(ok, what):
    print(("ok   " if ok else "FAIL ") + what, flush=True)
    if not ok:
        FAILED.append(what)
    return ok


def loop_checks():
    m = cpu.gen()
    for name in cpu.LOOP_CASES:
        icc, tree, kw = m.loop_case(name)
        z = np.load(os.path.join(cpu.GOLD, name + ".npz"))
        cells, tasks, nt = tree
        be = HydroBackend(parity.gpu_params(kw))
        be.build_from_cells(icc, cells, tasks, nt)
        # what a step leaves behind for the loop: the sorts, the cells' boxes and h_max (first density pass; the
        # smoothing lengths are converged already and stay as they are)
        be.runner_do_hydro_sort()
        be.runner_do_density()
        w = be.runner_do_limiter_loop(want_wakeup=True)
        nbad = int((w != z["l_wakeup"]).sum())
        check(nbad == 0, "%s: wakeup of %d particles (%d woken), bins on the device; %d differ" % (
            name, len(w), int((z["l_wakeup"] != il.NOT_AWAKE).sum()), nbad))
        # other bins handed over by a caller that runs the time-step task itself: everybody two bins up but the
        # starting ones -- then the same bins again
        up = np.where(icc["time_bin"] <= kw["max_active_bin"], icc["time_bin"], icc["time_bin"] + 2).astype(np.int8)
        w2 = be.runner_do_limiter_loop(time_bin=up, want_wakeup=True)
        want2 = m.run_loop_only("oracle", dict(icc, time_bin=up), tree, kw)
        check(np.array_equal(w2, want2), "%s: bins handed over (inactive ones two bins up) against the oracle" % name)
        w3 = be.runner_do_limiter_loop(time_bin=icc["time_bin"], want_wakeup=True)
        check(np.array_equal(w3, z["l_wakeup"]), "%s: the fixture's bins handed over again" % name)
        t = be.limiter_timings()
        check(t["limiter_loop_ms"] > 0.0, "%s: limiter loop device time %.3f ms" % (name, t["limiter_loop_ms"]))
        be.close()


def run_checks():
    m = cpu.gen()
    for name in cpu.RUN_CASES:
        icc, tree, kw, ip, variant, nsteps = m.run_case(name)
        z = np.load(os.path.join(cpu.GOLD, name + ".npz"))
        sphenix = variant == "sphenix"
        sx = ol.SPHENIX_PROPS if sphenix else None
        ref_states = []
        so, lo = il.run_loop("oracle", icc, kw, ip, tree, nsteps, variant=variant, sphenix=sx, limiter=True,
                             on_step=lambda step, s: ref_states.append({k: s[k].copy() for k in il.STATE_KEYS}))
        prm = parity.gpu_params(dict(kw, max_active_bin=il.NUM_TIME_BINS))
        be = HydroBackend(prm)
        if sphenix:
            be.sphenix_set_props(**ol.SPHENIX_PROPS)
        first_bad = []

        def on_step(step, g):
            bad = [k for k in il.STATE_KEYS
                   if not np.array_equal(np.asarray(g[k]).reshape(-1), ref_states[step][k].reshape(-1))]
            if bad and not first_bad:
                first_bad.append((step, bad))

        sg, lg = il.run_loop_gpu(be, icc, ip, tree, nsteps, on_step=on_step, sphenix=sphenix, limiter=True)
        t = be.limiter_timings()
        be.close()
        check(not first_bad, "%s: every field after every step against the oracle%s" % (
            name, "" if not first_bad else " -- first difference at step %d in %s" % first_bad[0]))
        check([l[:3] + l[5:] for l in lg] == [l[:3] + l[5:] for l in lo],
              "%s: per step active / sync point / woken, woken inactive, re-kicked: %s" % (name, [l[6:] for l in lg]))
        bad = [k for k in il.STATE_KEYS if not np.array_equal(np.asarray(sg[k]).reshape(-1), z[k].reshape(-1))]
        check(not bad, "%s: final state against the reference's fixture %s" % (name, bad))
        check([[l[0], l[1], l[2], l[5], l[6], l[7], l[8]] for l in lg] == z["log_int"].tolist(),
              "%s: log against the reference's fixture" % name)
        check(t["limiter_ms"] > 0.0 and t["limiter_loop_ms"] > 0.0, "%s: device times %s" % (name, t))


def refusal_checks():
    """What the entries refuse."""
    from swift_subtask_dev_b200 import ics
    ic = ics.uniform_random(12 ** 3, seed=5)
    be = HydroBackend(parity.gpu_params(dict(periodic=True)))
    try:
        be.runner_do_limiter_loop()
        check(False, "limiter loop on a context that is not built is refused")
    except B.SwiftCudaError:
        check(True, "limiter loop on a context that is not built is refused")
    be.build(ic)
    be.step()
    try:
        be.runner_do_limiter()
        check(False, "runner_do_limiter outside the time-integration mode is refused")
    except B.SwiftCudaError:
        check(True, "runner_do_limiter outside the time-integration mode is refused")
    # all bins equal: nobody can be woken
    w = be.runner_do_limiter_loop(want_wakeup=True)
    check((w == il.NOT_AWAKE).all(), "equal bins: nobody is woken")
    be.close()


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("loop", "all"):
        loop_checks()
    if what in ("run", "all"):
        run_checks()
    if what in ("refuse", "all"):
        refusal_checks()
    print("limiter_gpu_check:", "all ok" if not FAILED else "%d FAILED" % len(FAILED), flush=True)
    sys.exit(1 if FAILED else 0)
