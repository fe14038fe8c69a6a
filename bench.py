#!/usr/bin/env python
"""Benchmark of the SPH neighbour-loop hot path (density + ghost + force).

metric : ordered SPH pair interactions per second over one density+ghost+force
         step, whole job.
value  : device-timed, particles resident in HBM (CUDA events around the step).
e2e    : the same step through the reference-facing C-ABI entry with HOST buffers
         (swift_cuda_step_parts: the caller's array of 128-byte struct part records in, the same
         records out): H2D and D2H of the records inside the timed region.
--impl reference : the reference's own CPU operators (oracle/_ref, compiled from
         /root/reference; the plain-C oracle port when that is absent) on all
         host cores, on bounded samples of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "sph_pair_interactions_per_sec"
UNIT = "interactions/s"
FLOP_DENSITY, FLOP_FORCE = 66, 122  # algorithmic flop per ordered interaction


def workload(args):
    side = args.side
    return {"workload": "periodic uniform-random box %d^3 (%d particles), ~48 neighbours, density+ghost+force, "
                        "Minimal SPH + ideal gas" % (side, side ** 3),
            "side": side, "seed": 42,
            "l2": "inputs (>=180 B/particle resident) exceed the 126 MB L2 for side >= 96; no explicit flush",
            "sort_in_step": True,
            "step": "every timed step restarts from the initial h (device-side restore inside the timed region): "
                    "13-axis sort + density + all ghost redo sweeps + force"}


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nme, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------- CPU arm
def _cpu_worker(argtuple):
    side, seed, use_ref, use_fast = argtuple
    import oracle_lib as ol
    from swift_subtask_dev_b200 import ics
    ic = ics.uniform_random(side ** 3, seed=seed)
    prm = ol.make_params()
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    icc = ics.permute(ic, order)
    # the interactions are counted by the IEEE build (its recording wrapper); the TIMED run is the
    # reference built with its normal optimisation flags, which records nothing
    t0 = time.perf_counter()
    o = ol.run_step("ref" if use_ref else "oracle", prm, icc, cells, tasks, nt, record=False)
    dt = time.perf_counter() - t0
    if use_fast:
        t0 = time.perf_counter()
        ol.run_step("ref", prm, icc, cells, tasks, nt, record=False, variant="fast")
        dt = time.perf_counter() - t0
    return sum(o["counts"]), dt


def cpu_sample_text(nboxes, sample_side, fast):
    return ("%d independent %d^3 uniform-random periodic boxes (one per host process, %d^3 unsplit top-level cells "
            "each), sort + density + ghost + force by the reference's own runner_* operators compiled from "
            "/root/reference/src (%s), single thread each; NOT one %s box on the threaded engine"
            % (nboxes, sample_side, 12, "-O3 -ffast-math -funroll-loops -march=x86-64-v3, the reference's normal "
               "flags with a portable -march" if fast else "-O2 -fno-fast-math -ffp-contract=off", "256^3"))


def cpu_arm(sample_side, nproc, reps=1):
    """Reference CPU operators on `nproc` host processes, each stepping its own
    sample_side^3 sample of the workload.  Returns (interactions/s aggregate, description)."""
    import multiprocessing as mp
    import ctypes
    import oracle_lib as ol
    use_ref = ol.ref_available("minimal")
    use_fast = use_ref and ol.ref_available("fast")
    if not use_ref:
        ol.build_oracle()
    else:
        # the workers are forked children: load the libraries here too, so that whoever watches this
        # process sees which native code the arm runs
        ol.ref("minimal")
        if use_fast:
            ctypes.CDLL(os.path.join(ol.ORACLE_DIR, "_ref", "libswiftref_fast.so"))
    jobs = [(sample_side, 1000 + k, use_ref, use_fast) for k in range(nproc * reps)]
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(nproc) as pool:
        res = pool.map(_cpu_worker, jobs, chunksize=1)
    wall = time.perf_counter() - t0
    total = sum(r[0] for r in res)
    # tree building / IC generation are outside the operators: use the summed operator time per process
    busy = max(sum(r[1] for r in res[k::nproc]) for k in range(nproc))
    return total / busy, {"kind": "reference" if use_ref else "port", "cores": nproc,
                          "sample": cpu_sample_text(nproc * reps, sample_side, use_fast),
                          "build": "fast" if use_fast else ("ieee" if use_ref else "oracle port"),
                          "wall_s": wall, "interactions": total}


# ---- the threaded CPU arm: ONE box shared by all host threads ------------------------------------------------
_BOX = {}


def _box_setup(side):
    """Particles, the tree and task list, and the interaction count of one side^3 box (cached: the count comes from
    the reference built with IEEE flags and its pair-recording wrapper, single thread, untimed)."""
    if side not in _BOX:
        import oracle_lib as ol
        from swift_subtask_dev_b200 import ics
        ic = ics.uniform_random(side ** 3, seed=1000)
        prm = ol.make_params()
        order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
        icc = ics.permute(ic, order)
        ol.ref("minimal").ref_set_threads(1)
        o = ol.run_step("ref", prm, icc, cells, tasks, nt, record=False)
        _BOX[side] = (prm, icc, cells, tasks, nt, sum(o["counts"]))
    return _BOX[side]


def _threaded_child(side, nthreads, reps, q):
    try:
        import ctypes
        import oracle_lib as ol
        prm, icc, cells, tasks, nt, inter = _box_setup(side)
        lib = ol.ref("fast")
        lib.ref_set_threads.argtypes = [ctypes.c_int]
        lib.ref_last_step_seconds.restype = ctypes.c_double
        lib.ref_set_threads(nthreads)
        secs = []
        for _ in range(reps + 1):   # the first repetition warms the page cache and the allocator
            ol.run_step("ref", prm, icc, cells, tasks, nt, record=False, variant="fast")
            secs.append(lib.ref_last_step_seconds())
        q.put((inter, secs[1:]))
    except Exception as e:  # noqa: BLE001
        q.put(("error", repr(e)))


def cpu_arm_threaded(side, nthreads, reps=1, timeout_s=600):
    """The reference's own runner_* operators (compiled from /root/reference/src with its normal optimisation
    flags) on ONE side^3 box shared by `nthreads` pthreads -- oracle/ref_harness.c: ref_set_threads: tasks drawn
    from a shared counter under the locks of the top-level cells they update, the role of the reference's threaded
    runner (src/engine.c:1692-1736, src/runner_main.c, src/cell_lock.c).  Timed: the operator phases (sort,
    density, ghost, force, end_force) inside the harness; the tree and task build are outside, as for the GPU
    arm.  Runs in a child process so that a hang cannot take the bench with it.  Returns (interactions/s per
    repetition, description) or None when the compiled reference is not there or the child failed."""
    import multiprocessing as mp
    import oracle_lib as ol
    if not (ol.ref_available("minimal") and ol.ref_available("fast")):
        return None
    ctx = mp.get_context("fork")
    q = ctx.Queue()
    t0 = time.perf_counter()
    p = ctx.Process(target=_threaded_child, args=(side, nthreads, reps, q))
    p.start()
    try:
        inter, secs = q.get(timeout=timeout_s)
    except Exception:  # noqa: BLE001
        p.kill()
        return None
    p.join(10)
    if inter == "error":
        sys.stderr.write("threaded CPU arm failed: %s\n" % secs)
        return None
    vals = [inter / s for s in secs]
    return vals, {"kind": "reference", "cores": nthreads,
                  "sample": ("ONE %d^3 uniform-random periodic box (12^3 top-level cells%s), sort + density + ghost + force "
                             "by the reference's own runner_* operators compiled from /root/reference/src (-O3 -ffast-math "
                             "-funroll-loops -march=x86-64-v3, the reference's normal flags with a portable -march), shared by "
                             "%d threads: tasks drawn from a shared counter under the locks of the top-level cells they update "
                             "(oracle/ref_harness.c ref_set_threads; the reference's own scheduler and engine are not "
                             "built here); tree and task build outside the timed region"
                             % (side, ", each split once: sub-tasks" if side >= 96 else "", nthreads)),
                  "build": "fast", "wall_s": time.perf_counter() - t0, "interactions": inter,
                  "operator_s": secs}


def resident_leg(args, ic, hbm_peak):
    """the particles go to the device ONCE and stay there; per engine step only scalars
    (ti_end_min, dx_max) come back.  Step 0 has every particle active; the following ones the particles whose
    time bin ends (individual time steps), all particles drifted at every step (engine_policy_drift_all).
    Times are CUDA events on the library's stream.  DRIFT_BYTES: what k_drift reads + writes per particle."""
    from swift_subtask_dev_b200 import HydroBackend, make_params
    DRIFT_BYTES = 160 + 124
    n = len(ic["h"])
    ic = dict(ic)
    ic["time_bin"] = np.zeros(n, dtype=np.int8)   # hydro_first_init_part
    be = HydroBackend(make_params())
    be.build(ic)
    lmhc = float(np.log(np.float32(np.float32(1.4) ** np.float32(1.0 / 3.0)), dtype=np.float32))
    be.integrator_init(cfl_condition=0.1, log_max_h_change=lmhc, minimal_internal_energy=0.0, dt_min=1e-12,
                       dt_max=1e-2, time_base=1.0 / float(1 << 57))
    steps = []
    ti_current, ti_old = 0, 0
    for k in range(args.resident_steps + 1):
        mab = 56
        if ti_current:
            mab = 1
            while not ((1 << (mab + 1)) & ti_current):
                mab += 1
        be.set_time(ti_current, mab)
        if k > 0:
            be.drift(ti_old)
        be.step(do_sort=True)
        cnt, t = be.counters(), be.timings()
        if k > 0:
            be.kick(2)
        ti_end_min, updated = be.timestep()
        be.kick(1)
        it = be.integrator_timings()
        integ_ms = (it["drift_ms"] if k > 0 else 0.0) + (it["kick2_ms"] if k > 0 else 0.0) + it["kick1_ms"] + \
            it["timestep_ms"]
        steps.append({"active": updated, "interactions": cnt["n_density"] + cnt["n_density_redo"] + cnt["n_force"],
                      "loops_ms": t["total_ms"], "integrator_ms": integ_ms,
                      "phase_ms": [round(t[q], 3) for q in ("sort_ms", "density_ms", "ghost_ms", "force_ms")],
                      "drift_gbs": DRIFT_BYTES * n / (it["drift_ms"] * 1e-3) / 1e9 if k > 0 else None})
        ti_old, ti_current = ti_current, ti_end_min
    be.close()
    later = steps[1:]
    tot_ms = sum(s["loops_ms"] + s["integrator_ms"] for s in steps)
    out = {"steps": steps, "value": sum(s["interactions"] for s in steps) / (tot_ms * 1e-3), "unit": UNIT,
           "host_bytes_per_step": 24, "note": "whole engine steps with the particles resident in HBM; value = "
           "interactions of all steps / device time of all steps (loops + drift + kicks + time-step)"}
    if later:
        g = [s["drift_gbs"] for s in later]
        out["drift_kernel"] = {"bound": "hbm", "achieved": float(np.mean(g)), "peak": hbm_peak, "unit": "GB/s",
                               "frac": float(np.mean(g)) / hbm_peak, "bytes_per_particle": DRIFT_BYTES}
    return out


def sphenix_leg(side):
    """one step of the three-loop scheme (sort, density, ghost, GRADIENT loop, extra ghost, force)
    on a side^3 uniform-random periodic box, CUDA events of the library, in both fields modes (EXACT: every field
    bit-identical to the reference compiled with SPHENIX_SPH; FAST: the force interaction with FMA / MUFU arithmetic,
    within 1e-4).  Repetitions restart from the initial h; the viscosity / diffusion alphas carry over, as in a run."""
    from swift_subtask_dev_b200 import HydroBackend, ics, make_params
    from swift_subtask_dev_b200 import backend as B
    ic = ics.uniform_random(side ** 3, seed=42)
    out = {"workload": "uniform-random periodic box %d^3, SPHENIX + ideal gas" % side,
           "note": "ghost_ms includes the gradient loop and the extra ghost"}
    for mode, name in ((1, "fast"), (0, "exact")):
        be = HydroBackend(make_params(periodic=True, alpha=0.1, scheme=B.SCHEME_SPHENIX, fields_mode=mode))
        be.build(ic)
        best = None
        for rep in range(4):  # the first one is the warm-up
            be.step(do_sort=True, restore=True)
            t = be.timings()
            if rep > 0 and (best is None or t["total_ms"] < best["total_ms"]):
                best = t
        c = be.counters()
        ngrad = be.sphenix_results()["n_gradient"]
        be.close()
        inter = c["n_density"] + c["n_density_redo"] + c["n_force"] + ngrad
        out[name] = {"ms_per_step": best["total_ms"], "phase_ms": {k: round(v, 3) for k, v in best.items()},
                     "n_gradient": ngrad, "interactions_per_step": inter, "value": inter / (best["total_ms"] * 1e-3),
                     "unit": UNIT}
    return out


def limiter_leg(side, timeout_s=150):
    """The time-step limiter's device times on a resident run (scripts/limiter_bench.py), in a process of its own:
    its kernels were written after the last GPU session of the round, so nothing they do -- a refusal, a device
    fault, a hang -- may touch this process or its JSON line.  Returns the script's line, or why there is none."""
    import subprocess
    try:
        r = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "scripts",
                                                         "limiter_bench.py"), str(side)],
                           capture_output=True, text=True, timeout=timeout_s)
        if r.returncode != 0:
            return {"unavailable": "limiter_bench.py exited %d: %s" % (r.returncode, (r.stderr or r.stdout)[-300:])}
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001  (a timeout, no output, bad JSON: the bench line goes out regardless)
        return {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# --------------------------------------------------------------------------- main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--side", type=int, default=256, help="particles per box side (BASELINE configs[1]: 256)")
    ap.add_argument("--cpu-side", type=int, default=64, help="side of the independent boxes of the fallback CPU arm")
    ap.add_argument("--cpu-box-side", type=int, default=128,
                    help="side of the ONE box the threaded CPU arm steps with all host threads (0: the independent-boxes arm)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--workload", default="uniform", choices=["uniform", "clustered"],
                    help="uniform: BASELINE configs[1] / [3] (default, the metric's config); clustered: configs[4] "
                         "shape, side^3 particles in TOTAL over ~10^3 clumps + background")
    ap.add_argument("--fields", default="fast", choices=["fast", "exact"],
                    help="swift_cuda_params.fields_mode of the headline run: fast = FMA / MUFU arithmetic in the force "
                         "loop (pair lists, counts and h bit-exact, a / u_dt / h_dt / v_sig within 1e-4); exact = every "
                         "field bit-identical to the reference.  The other mode is timed too (detail.other_mode)")
    ap.add_argument("--sphenix-side", type=int, default=128,
                    help="box side of the SPHENIX leg (three-loop scheme, SURVEY 8(f) N1; detail.sphenix); 0: skip")
    ap.add_argument("--limiter-side", type=int, default=96,
                    help="lattice side of the time-step limiter leg (resident engine steps with the limiter on, in a "
                         "process of its own; detail.limiter); 0: skip")
    ap.add_argument("--resident-steps", type=int, default=3,
                    help="engine steps of the device-resident leg (drift/kick/time-step kernels, SURVEY 8(f) N3); 0: skip")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = workload(args)

    if args.impl == "reference":
        if rank != 0:
            return 0
        nproc = host_cores()
        t0 = time.perf_counter()
        thr = cpu_arm_threaded(args.cpu_box_side, nproc, reps=max(args.steps, 1)) if args.cpu_box_side > 0 else None
        if thr is not None:
            vals, desc = thr
            dt = float(np.sum(desc["operator_s"]))
        else:
            for _ in range(max(args.warmup, 0) and 1):
                cpu_arm(32, nproc)
            t0 = time.perf_counter()
            vals = []
            for _ in range(args.steps):
                v, desc = cpu_arm(args.cpu_side, nproc)
                vals.append(v)
            dt = time.perf_counter() - t0
        v = float(np.mean(vals))
        desc["value"] = v
        desc["unit"] = UNIT
        # what this arm ran is NOT the 256^3 box of the b200 arm: say so in its config
        cfg = dict(cfg)
        cfg["workload"] = "CPU sample of the b200 arm's workload: " + desc["sample"]
        cfg.pop("step", None)
        cfg.pop("l2", None)
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": cfg, "cpu_baseline": desc,
                          "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return 0

    import torch
    from swift_subtask_dev_b200 import HydroBackend, ics, make_params
    if world > 1:
        from swift_subtask_dev_b200 import multi
        return multi.bench_multi(args, cfg, METRIC, UNIT, sampler_cls=ClockSampler)

    torch.cuda.set_device(0)
    if args.workload == "clustered":
        ic = ics.clustered(args.side ** 3, nclumps=1000, seed=7)
        cfg["workload"] = ("clustered periodic box (1000 Plummer clumps + 30 %% uniform background), %d^3 particles, "
                           "density+ghost+force, Minimal SPH + ideal gas" % args.side)
    else:
        ic = ics.uniform_random(args.side ** 3, seed=cfg["seed"])
    n = len(ic["h"])
    fast = args.fields == "fast"
    cfg["fields_mode"] = ("FAST: neighbour lists, counts and h bit-exact; force-loop fields (a, u_dt, h_dt, v_sig) with FMA "
                          "/ MUFU arithmetic, within 1e-4 of the reference" if fast else
                          "EXACT: every field bit-identical to the reference")
    # the other mode first, briefly: its step time goes into detail.other_mode
    be = HydroBackend(make_params(fields_mode=0 if fast else 1))
    be.build(ic)
    for _ in range(2):
        be.step(restore=True)
    other_ms = be.timings()["total_ms"]
    other_cnt = be.counters()
    be.close()
    be = HydroBackend(make_params(fields_mode=1 if fast else 0))
    t0 = time.perf_counter()
    be.build(ic)
    build_s = time.perf_counter() - t0

    # every step starts from the SAME state (the h of the initial conditions is put back on the
    # device inside the timed region), so each timed step does the work of the first one:
    # sort, density, all ghost redo sweeps, force.
    for _ in range(args.warmup):
        be.step(restore=True)
    sampler = ClockSampler(0)
    sampler.start()
    torch.cuda.synchronize()
    times, phase = [], {"sort_ms": 0.0, "density_ms": 0.0, "ghost_ms": 0.0, "force_ms": 0.0}
    launches = 0
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        be.step(restore=True)
        t = be.timings()
        times.append(t["total_ms"])
        for k in phase:
            phase[k] += t[k]
        launches += be.counters()["n_kernel_launches"]
    torch.cuda.synchronize()
    wall_ms = 1e3 * (time.perf_counter() - wall0) / args.steps
    clocks = sampler.stop()
    cnt = be.counters()
    inter = cnt["n_density"] + cnt["n_density_redo"] + cnt["n_force"]
    ms = float(np.mean(times))
    value = inter / (ms * 1e-3)

    # ---- end to end through the reference-facing entry: the caller's array of 128-byte struct part
    # records (pinned host memory) in, the same records out (swift_cuda_step_parts: H2D of the records,
    # device-side unpack to the SoA caches, the step, device-side write-back into the records, D2H)
    from swift_subtask_dev_b200 import backend as be_mod
    # The records are laid out as s->parts is after space_rebuild / space_split (every cell's particles
    # contiguous: a struct cell's hydro.parts is a pointer into that array), so the entry can hand the force
    # loop's results back in pieces, the copy of one piece overlapping the loop over the next.
    cell_order = be.tree()["order"]
    ic_cells = ics.permute(ic, cell_order)
    parts = be_mod.parts_from_ic(ic_cells, pinned=True)
    bp = HydroBackend(make_params(fields_mode=1 if fast else 0))
    be.close()
    bp.build_from_parts(parts)
    e2e_times = []
    for k in range(max(1, min(args.steps, 3)) + 1):
        parts["h"] = ic_cells["h"]             # every step from the initial h (host side, outside the timing)
        t0 = time.perf_counter()
        bp.step_parts(parts)
        dt = time.perf_counter() - t0
        if k > 0:
            e2e_times.append(dt)
    h2d = d2h = n * parts.dtype.itemsize
    cnt_p = bp.counters()
    assert cnt_p["n_density"] + cnt_p["n_density_redo"] + cnt_p["n_force"] == inter
    e2e_value = inter / float(np.mean(e2e_times))
    bp.close()
    be = HydroBackend(make_params(fields_mode=1 if fast else 0))
    be.build(ic)
    # ---- the same through the SoA entry (swift_cuda_step_io: five field arrays in, result arrays out,
    # transfers overlapped with the kernels): reported as detail.e2e_soa
    soa_times = []
    h2d_soa = n * (24 + 12 + 4 + 4 + 4)        # x, v, mass, h, u from pinned host memory
    d2h_soa = n * (4 * (6 + 3 + 3) + 1)        # h, rho, P, c, f, balsara, a[3], u_dt, h_dt, v_sig, min_ngb_time_bin
    ic_pin = {}
    keep = []
    for k, a in ic.items():
        if isinstance(a, np.ndarray):
            t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
            keep.append(t)
            ic_pin[k] = t.numpy()
        else:
            ic_pin[k] = a
    out_pin = be.alloc_results(pinned=True)
    for k in range(max(1, min(args.steps, 2)) + 1):
        t0 = time.perf_counter()
        be.step_io(ic_pin, out_pin)   # upload + sort + density + ghost + force + download, one C-ABI call
        dt = time.perf_counter() - t0
        if k > 0:
            soa_times.append(dt)
    e2e_soa = {"value": inter / float(np.mean(soa_times)), "unit": UNIT, "h2d_bytes_per_step": h2d_soa,
               "d2h_bytes_per_step": d2h_soa, "entry": "swift_cuda_step_io (SoA field arrays)"}

    # ---- roofline of the dominant kernel: the single launch that takes longest (the force loop; the ghost
    # phase is longer in total but is a dozen launches of k_density<1> / k_density_list).  FP32 CUDA-core bound (SURVEY.md
    # 8(d)): algorithmic flops = 66 per density interaction, 122 per force interaction; candidate tests
    # that miss are not credited.  kernel_ms = CUDA events on the library's stream around that launch.
    fname = "k_force_fast" if fast else "k_force"
    per_launch = {"k_density<0>": (phase["density_ms"] / args.steps, FLOP_DENSITY * cnt["n_density"]),
                  fname: (phase["force_ms"] / args.steps, FLOP_FORCE * cnt["n_force"])}
    dom = max(per_launch, key=lambda k: per_launch[k][0])
    kern_ms, flops = per_launch[dom]
    achieved = flops / (kern_ms * 1e-3) / 1e12
    fp32_measured = be.measure_fp32_peak()
    peak_nominal = 148 * 128 * 2 * 1965e6 / 1e12
    peak = fp32_measured if fp32_measured > 0 else peak_nominal
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6400.0))
    # dram bytes of that kernel per launch, from the committed `ncu --set full` capture of this command
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            tj = json.load(f)
        if tj.get("side") == args.side:
            traffic = tj["dram_bytes_per_launch"].get(dom)
    except Exception:
        pass
    redo_ms = phase["ghost_ms"] / args.steps
    roofline = {"bound": "fp32", "kernel": dom,
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "peak_source": "FALLBACK denominator: FMA microbenchmark run by this bench on this GPU at the clocks in "
                               "`clocks` (MEASURED_PEAKS.json has no FP32 entry); nominal 148 SMs x 128 lanes x 2 x "
                               "1965 MHz = %.1f" % peak_nominal,
                "peak_sm_mhz": clocks.get("sm_mhz"),
                "kernel_ms": kern_ms, "traffic": traffic,
                "note": "bound is the FP32 CUDA-core issue rate (SURVEY 8(d): ~40 flop/B, no dense contraction: "
                        "neither 'hbm' nor 'tensor'); achieved counts only the 122 (66) algorithmic flop of "
                        "interactions that hit",
                "other_kernels": {"k_density<0>": {"ms": per_launch["k_density<0>"][0],
                                                   "tflops": per_launch["k_density<0>"][1] /
                                                   (per_launch["k_density<0>"][0] * 1e-3) / 1e12},
                                  "ghost: k_density<1> + k_density_list x %d sweeps + k_ghost" % cnt["ghost_sweeps"]:
                                      {"ms": redo_ms,
                                       "tflops": FLOP_DENSITY * cnt["n_density_redo"] / (redo_ms * 1e-3) / 1e12}},
                "hbm_step": {"achieved_gbs": 223.0 * n / (ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                             "frac": 223.0 * n / (ms * 1e-3) / 1e9 / hbm_peak,
                             "note": "223 B/particle/step algorithmic (SURVEY 8(d)) over the whole step; "
                                     "peak = MEASURED_PEAKS.json hbm_gbs"}}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "entry": "swift_cuda_step_parts (the caller's 128-byte struct part records in and out; records in cell order as s->parts is after space_rebuild, taken in 12 pieces: unpack, sorts and the first density pass of a piece run under the copy of the next, results are copied back under the force loop)"},
            "roofline": roofline,
            "detail": {"n_particles": n, "interactions_per_step": inter, "n_density": cnt["n_density"],
                       "n_density_redo": cnt["n_density_redo"], "n_force": cnt["n_force"],
                       "ghost_sweeps": cnt["ghost_sweeps"], "phase_ms": {k: v / args.steps for k, v in phase.items()},
                       "wall_ms_per_step": wall_ms, "build_s": build_s, "leaf_interactions": cnt["n_leaf_interactions"],
                       "e2e_soa": e2e_soa,
                       "other_mode": {"fields_mode": "EXACT" if fast else "FAST", "ms_per_step": other_ms,
                                      "value": inter / (other_ms * 1e-3),
                                      "same_counts": bool(other_cnt["n_density"] == cnt["n_density"] and
                                                          other_cnt["n_density_redo"] == cnt["n_density_redo"] and
                                                          other_cnt["n_force"] == cnt["n_force"])}}}
    be.close()
    if args.resident_steps > 0:
        line["resident"] = resident_leg(args, ic, hbm_peak)
    if args.sphenix_side > 0:
        line["detail"]["sphenix"] = sphenix_leg(args.sphenix_side)
    if args.limiter_side > 0:
        line["detail"]["limiter"] = limiter_leg(args.limiter_side)
    if not args.no_cpu:
        nproc = host_cores()
        thr = cpu_arm_threaded(args.cpu_box_side, nproc, reps=2) if args.cpu_box_side > 0 else None
        if thr is not None:
            v, desc = float(np.mean(thr[0])), thr[1]
            # the arm of round 1 beside it: one independent single-thread box per core (no shared box, no locks)
            vi, di = cpu_arm(args.cpu_side, nproc)
            desc["independent_boxes"] = {"value": vi, "cores": nproc, "sample": di["sample"]}
        else:
            v, desc = cpu_arm(args.cpu_side, nproc)
        desc["value"] = v
        desc["unit"] = UNIT
        line["cpu_baseline"] = desc
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    sys.exit(main())
