"""Device time of the time-step limiter on a resident run: python scripts/limiter_bench.py [side] [steps]
A side^3 lattice with a Sedov-like energy spike (ics.lattice), particles resident on the device, engine steps with
individual time bins and the limiter on (drift, hydro loops, kick2, time-step, limiter loop, runner_do_limiter, kick1).
Prints ONE JSON line: per step the active and woken particles and the CUDA-event times of the limiter loop
(k_density<0,false,2>) and of k_limiter.  bench.py runs it in a process of its own (detail.limiter); parity of the
two kernels is the business of tests/test_zyz_gpu_limiter.py, not of this script."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from swift_subtask_dev_b200 import HydroBackend, ics, make_params  # noqa: E402


def main():
    side = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    ic = ics.lattice(side, jitter=0.05, sedov=True)
    n = len(ic["h"])
    ic["time_bin"] = np.zeros(n, dtype=np.int8)
    be = HydroBackend(make_params())
    be.build(ic)
    lmhc = float(np.log(np.float32(np.float32(1.4) ** np.float32(1.0 / 3.0)), dtype=np.float32))
    be.integrator_init(cfl_condition=0.1, log_max_h_change=lmhc, minimal_internal_energy=0.0, dt_min=1e-14,
                       dt_max=1e-2, time_base=1.0 / float(1 << 57))
    steps = []
    ti_current, ti_old = 0, 0
    for k in range(nsteps + 1):
        mab = 56
        if ti_current:
            mab = 1
            while not ((1 << (mab + 1)) & ti_current):
                mab += 1
        be.set_time(ti_current, mab)
        if k > 0:
            be.drift(ti_old)
        be.step(do_sort=True)
        loops_ms = be.timings()["total_ms"]
        if k > 0:
            be.kick(2)
        ti_end_min, updated = be.timestep()
        be.runner_do_limiter_loop()
        t_lim, woken = be.runner_do_limiter()
        ti_end_min = min(ti_end_min, t_lim)
        be.kick(1)
        lt = be.limiter_timings()
        steps.append({"active": int(updated), "woken": int(woken), "loops_ms": round(loops_ms, 3),
                      "limiter_loop_ms": round(lt["limiter_loop_ms"], 4), "limiter_ms": round(lt["limiter_ms"], 4)})
        ti_old, ti_current = ti_current, ti_end_min
    be.close()
    print(json.dumps({"workload": "%d^3 lattice, Sedov-like spike, resident engine steps with the limiter" % side,
                      "n_particles": n, "steps": steps,
                      "note": "limiter_loop_ms: k_density<0,false,2> (the first-pass walk once more, scattering wake-up "
                              "flags); limiter_ms: k_limiter (runner_do_limiter); loops_ms: sort + density + ghost + "
                              "force of the same step"}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
