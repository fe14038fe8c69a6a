"""TEST-ONLY bindings of the time-integration checkers: oracle_drift / oracle_kick /
oracle_timestep of oracle/sph_oracle.c and ref_drift / ref_kick / ref_timestep of oracle/ref_harness.c (the
reference's own cell_drift_part, runner_do_kick1/2, runner_do_timestep on one cell holding all particles),
and the engine-step loop both are driven through:

    step 0:  all active: sort, density, ghost, force; timestep; kick1
    step k:  ti_current = ti_end_min; max_active_bin = get_max_active_bin(ti_current); drift ALL particles
             (the reference's engine_policy_drift_all); sort, density, ghost, force on the active ones;
             kick2; timestep; kick1                                    (src/engine.c engine_step order)
"""
import ctypes as C

import numpy as np

import oracle_lib as ol

NUM_TIME_BINS = 56
_FP = C.POINTER(C.c_float)
_BP = C.POINTER(C.c_byte)


class IParams(C.Structure):
    _fields_ = [("scheme", C.c_int), ("cfl_condition", C.c_float), ("log_max_h_change", C.c_float),
                ("minimal_internal_energy", C.c_float), ("h_max", C.c_float), ("h_min", C.c_float),
                ("dt_min", C.c_double), ("dt_max", C.c_double), ("time_base", C.c_double),
                ("max_active_bin", C.c_int), ("ti_current", C.c_longlong)]


class IState(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.POINTER(C.c_double)), ("v", _FP), ("v_full", _FP), ("u", _FP),
                ("u_full", _FP), ("h", _FP), ("rho", _FP), ("pressure", _FP), ("soundspeed", _FP), ("v_sig", _FP),
                ("a", _FP), ("u_dt", _FP), ("h_dt", _FP), ("x_diff", _FP), ("x_diff_sort", _FP),
                ("time_bin", _BP), ("min_ngb_time_bin", _BP), ("mat_id", C.POINTER(C.c_int))]


STATE_F32 = ("v", "v_full", "u", "u_full", "h", "rho", "pressure", "soundspeed", "v_sig", "a", "u_dt", "h_dt",
             "x_diff", "x_diff_sort")
STATE_KEYS = ("x",) + STATE_F32 + ("time_bin",)


def default_iparams(scheme=0, cfl=0.1, h_max=1e10, h_min=0.0, dt_min=1e-12, dt_max=1e-2, time_end=1.0,
                    min_u=0.0):
    """hydro_props_init defaults: log_max_h_change = logf(powf(1.4, 1/3)) (src/hydro_properties.c:122, 364);
    time_base = (time_end - time_begin) / max_nr_timesteps (src/engine.c)."""
    return dict(scheme=scheme, cfl_condition=cfl,
                log_max_h_change=float(np.log(np.float32(np.float32(1.4) ** np.float32(1.0 / 3.0)), dtype=np.float32)),
                minimal_internal_energy=min_u, h_max=h_max, h_min=h_min, dt_min=dt_min, dt_max=dt_max,
                time_base=time_end / float(1 << (NUM_TIME_BINS + 1)))


def get_max_active_bin(ti):
    """src/timeline.h:146-154"""
    if ti == 0:
        return NUM_TIME_BINS
    b = 1
    while not ((1 << (b + 1)) & ti):
        b += 1
    return b


def new_state(ic):
    """hydro_first_init_part (Minimal/hydro.h:983-995): v_full = v, u_full = u; everything else zero."""
    n = len(ic["h"])
    s = {"x": np.array(ic["x"], dtype=np.float64, order="C"), "v": np.array(ic["v"], dtype=np.float32, order="C"),
         "u": np.array(ic["u"], dtype=np.float32), "h": np.array(ic["h"], dtype=np.float32)}
    s["v_full"] = s["v"].copy()
    s["u_full"] = s["u"].copy()
    for k in ("rho", "pressure", "soundspeed", "v_sig", "u_dt", "h_dt"):
        s[k] = np.zeros(n, dtype=np.float32)
    for k in ("a", "x_diff", "x_diff_sort"):
        s[k] = np.zeros((n, 3), dtype=np.float32)
    s["time_bin"] = np.zeros(n, dtype=np.int8)   # hydro_first_init_part: bin 0, active in step 0
    s["min_ngb_time_bin"] = np.full(n, NUM_TIME_BINS + 1, dtype=np.int8)
    s["mat_id"] = np.array(ic["mat_id"], dtype=np.int32) if ic.get("mat_id") is not None else None
    return s


def _cstate(s):
    c = IState()
    c.n = len(s["h"])
    c.x = s["x"].ctypes.data_as(C.POINTER(C.c_double))
    for k in STATE_F32:
        assert s[k].dtype == np.float32 and s[k].flags["C_CONTIGUOUS"], k
        setattr(c, k, s[k].ctypes.data_as(_FP))
    c.time_bin = s["time_bin"].ctypes.data_as(_BP)
    c.min_ngb_time_bin = s["min_ngb_time_bin"].ctypes.data_as(_BP)
    if s.get("mat_id") is not None:
        c.mat_id = s["mat_id"].ctypes.data_as(C.POINTER(C.c_int))
    return c


def _ciparams(ip, ti_current, max_active_bin):
    c = IParams()
    for k, v in ip.items():
        setattr(c, k, v)
    c.ti_current = ti_current
    c.max_active_bin = max_active_bin
    return c


class Checker:
    """which = "oracle" (C restatement) or "ref" (compiled reference)."""

    def __init__(self, which, ip, eos=None, variant="minimal"):
        self.which, self.ip, self.variant = which, dict(ip), variant
        self.eos = eos if eos is not None else ol.Eos()
        self.lib = ol.oracle() if which == "oracle" else ol.ref(variant)
        self.pre = "oracle_" if which == "oracle" else "ref_"

    def _call(self, name, cip, s, *extra):
        fn = getattr(self.lib, self.pre + name)
        fn.restype = C.c_int
        args = [C.byref(cip)]
        if self.which == "oracle" and name in ("drift", "kick"):
            args.append(C.byref(self.eos))
        args.append(C.byref(_cstate(s)))
        return fn(*args, *extra)

    def drift(self, s, ti_old, ti_current):
        dx = (C.c_float * 2)()
        rc = self._call("drift", _ciparams(self.ip, ti_current, -1), s, C.c_longlong(ti_old), dx)
        assert rc == 0
        return float(dx[0]), float(dx[1])

    def kick(self, s, which, ti_current, max_active_bin):
        rc = self._call("kick", _ciparams(self.ip, ti_current, max_active_bin), s, C.c_int(which))
        assert rc == 0

    def timestep(self, s, ti_current, max_active_bin):
        ti_end_min, updated = C.c_longlong(0), C.c_longlong(0)
        rc = self._call("timestep", _ciparams(self.ip, ti_current, max_active_bin), s, C.byref(ti_end_min),
                        C.byref(updated))
        assert rc == 0, "a particle wants a time-step below dt_min"
        return ti_end_min.value, updated.value

    def limiter(self, s, wakeup, ti_current, max_active_bin):
        """runner_do_limiter: timestep_limit_part for every particle with wakeup != -56; wakeup is reset in place.
        Returns (earliest new end of step or max_nr_timesteps, number of limited particles)."""
        assert wakeup.dtype == np.int8 and wakeup.flags["C_CONTIGUOUS"]
        ti_end_min, limited = C.c_longlong(0), C.c_longlong(0)
        fn = getattr(self.lib, self.pre + "limiter")
        fn.restype = C.c_int
        args = [C.byref(_ciparams(self.ip, ti_current, max_active_bin))]
        if self.which == "oracle":
            args.append(C.byref(self.eos))
        rc = fn(*args, C.byref(_cstate(s)), wakeup.ctypes.data_as(_BP), C.byref(ti_end_min), C.byref(limited))
        assert rc == 0
        return ti_end_min.value, limited.value


SPHENIX_STATE = (("alpha_visc", "e_alpha_visc"), ("alpha_diff", "e_alpha_diff"), ("div_v_prev", "e_div_v_prev"),
                 ("div_v_dt", "e_div_v_dt"), ("laplace_u", "e_laplace_u"))


def hydro_step(which, s, prm_kwargs, tree, max_active_bin, eos=None, variant="minimal", mass=None, ids=None,
               h_build=None, sphenix=None):
    """sort + density + ghost + force of the ACTIVE particles through the step checker; the inactive ones keep
    their state (their rho, P, f, c, balsara are what the force loop of the active ones reads)."""
    cells, tasks, ntasks = tree
    kw = dict(prm_kwargs)
    kw["max_active_bin"] = max_active_bin
    prm = ol.make_params(**kw)
    ic = {"x": s["x"], "v": s["v"], "mass": mass, "h": s["h"], "u": s["u"], "id": ids, "time_bin": s["time_bin"],
          "in_rho": s["rho"], "in_pressure": s["pressure"], "in_f": s["f"], "in_soundspeed": s["soundspeed"],
          "in_balsara": s["balsara"], "in_h_build": h_build}
    if s.get("mat_id") is not None:
        ic["mat_id"] = s["mat_id"]
    if sphenix is not None:
        # the three-loop scheme: the viscosity / diffusion state travels in the particle from step to step
        ic["in_alpha_visc"], ic["in_alpha_diff"], ic["in_div_v_prev"] = s["alpha_visc"], s["alpha_diff"], s["div_v_prev"]
    o = ol.run_step(which, prm, ic, cells, tasks, ntasks, record=False, variant=variant, eos=eos, sphenix=sphenix)
    act = s["time_bin"] <= max_active_bin
    if sphenix is not None:
        for ks, ko in SPHENIX_STATE:
            s[ks][act] = o[ko][act]
    for ks, ko in (("h", "g_h"), ("rho", "g_rho"), ("pressure", "g_pressure"), ("soundspeed", "g_soundspeed"),
                   ("f", "g_f"), ("balsara", "g_balsara"), ("a", "f_a"), ("u_dt", "f_u_dt"), ("h_dt", "f_h_dt"),
                   ("v_sig", "f_v_sig")):
        s[ks][act] = np.asarray(o[ko], dtype=np.float32).reshape(s[ks].shape)[act]
    s["min_ngb_time_bin"][act] = np.asarray(o["f_min_ngb_time_bin"], dtype=np.int8)[act]
    return o


NOT_AWAKE = -NUM_TIME_BINS   # time_bin_not_awake, src/timeline.h:48


def limiter_loop(which, s, prm_kwargs, tree, max_active_bin, variant="minimal", mass=None, ids=None, h_build=None,
                 sphenix=None):
    """The tasks of subtype limiter (runner_do{self,pair}1_branch_limiter, runner_dosub_*1_limiter) on the state
    after the time-step task: p->limiter_data.wakeup of every particle."""
    cells, tasks, ntasks = tree
    kw = dict(prm_kwargs)
    kw["max_active_bin"] = max_active_bin
    prm = ol.make_params(**kw)
    ic = {"x": s["x"], "v": s["v"], "mass": mass, "h": s["h"], "u": s["u"], "id": ids, "time_bin": s["time_bin"],
          "in_h_build": h_build, "l_time_bin": s["time_bin"], "l_only": True}
    if s.get("mat_id") is not None:
        ic["mat_id"] = s["mat_id"]
    o = ol.run_step(which, prm, ic, cells, tasks, ntasks, record=False, variant=variant, sphenix=sphenix)
    return o["l_wakeup"]


def limiter_census(wake, time_bin, max_active_bin):
    """(woken particles, of them inactive ones -- kick1 undone, src/timestep_limiter.h:87-215 --, of those the ones
    whose new bin does not start now and that get their new kick1 on the spot, :188-207)."""
    woken = wake != NOT_AWAKE
    inactive = woken & (time_bin > max_active_bin)
    rekicked = inactive & ((2 - wake.astype(np.int32)) > max_active_bin)
    return int(woken.sum()), int(inactive.sum()), int(rekicked.sum())


def run_loop(which, ic, prm_kwargs, ip, tree, nsteps, eos=None, variant="minimal", on_step=None, sphenix=None,
             limiter=False):
    """The engine-step loop of the module docstring on the checker `which`.  ic is in CELL order (tree built on
    it).  Returns the state dict and the per-step log [(ti_current, max_active_bin, updated, dx_max...)].
    limiter: the time-step limiter (swift --limiter) -- after the time-step task the limiter loop, then
    runner_do_limiter, then kick1 (engine_maketasks.c: timestep -> limiter tasks -> timestep_limiter -> kick1);
    the log entries gain limiter_census."""
    s = new_state(ic)
    n = len(s["h"])
    s["f"] = np.zeros(n, dtype=np.float32)
    s["balsara"] = np.zeros(n, dtype=np.float32)
    if sphenix is not None:
        # hydro_convert_quantities: the initial alphas of hydro_props; e->time_base is the integrator's
        sphenix = dict(sphenix, time_base=ip["time_base"])
        s["alpha_visc"] = np.full(n, prm_kwargs.get("alpha", 0.8), dtype=np.float32)
        s["alpha_diff"] = np.full(n, sphenix.get("diffusion_alpha", 0.0), dtype=np.float32)
        for k in ("div_v_prev", "div_v_dt", "laplace_u"):
            s[k] = np.zeros(n, dtype=np.float32)
    mass = np.array(ic["mass"], dtype=np.float32)
    ids = np.array(ic["id"], dtype=np.int64) if "id" in ic else np.arange(n, dtype=np.int64)
    chk = Checker(which, ip, eos=eos, variant=variant)
    h_build = s["h"].copy()   # c->hydro.h_max_old stays what it was at the rebuild
    log = []
    ti_current, ti_old = 0, 0
    for step in range(nsteps + 1):
        mab = get_max_active_bin(ti_current)
        dx = (0.0, 0.0)
        if step > 0:
            dx = chk.drift(s, ti_old, ti_current)
        hydro_step(which, s, prm_kwargs, tree, mab, eos=eos, variant=variant, mass=mass, ids=ids, h_build=h_build,
                   sphenix=sphenix)
        if step > 0:
            chk.kick(s, 2, ti_current, mab)
        ti_end_min, updated = chk.timestep(s, ti_current, mab)
        if limiter:
            wake = limiter_loop(which, s, prm_kwargs, tree, mab, variant=variant, mass=mass, ids=ids, h_build=h_build,
                                sphenix=sphenix)
            lim = limiter_census(wake, s["time_bin"], mab)
            t_lim, nlim = chk.limiter(s, wake, ti_current, mab)
            assert nlim == lim[0]
            ti_end_min = min(ti_end_min, t_lim)
        chk.kick(s, 1, ti_current, mab)
        log.append((ti_current, mab, updated, dx[0], dx[1], ti_end_min) + (lim if limiter else ()))
        if on_step is not None:
            on_step(step, s)
        ti_old, ti_current = ti_current, ti_end_min
    return s, log


def prepare(ic, prm_kwargs):
    """Converged smoothing lengths first (the reference's engine_init_particles does a "fake" step and
    rebuilds before the first real one), then the tree and task list on them.  Returns the particles in
    cell order and (cells, tasks, ntasks)."""
    prm = ol.make_params(**prm_kwargs)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    from swift_subtask_dev_b200 import ics
    icc = ics.permute(ic, order)
    eos = ol.eos_from_json() if prm_kwargs.get("scheme", 0) == 1 else None
    o0 = ol.run_step("oracle", prm, icc, cells, tasks, nt, eos=eos,
                     sphenix=ol.SPHENIX_PROPS if prm_kwargs.get("scheme", 0) == 2 else None)
    ic = dict(ic)
    h = np.empty_like(np.asarray(ic["h"], dtype=np.float32))
    h[order] = o0["g_h"]
    ic["h"] = h
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    return ics.permute(ic, order), (cells, tasks, nt)


def run_loop_gpu(be, ic, ip, tree, nsteps, on_step=None, sphenix=False, limiter=False):
    """The same loop through the C ABI: the particles go up once (build_from_cells) and stay on the
    device; per step only scalars (ti_end_min, dx_max) come back.  `be` is a HydroBackend created with
    max_active_bin = 56."""
    cells, tasks, nt = tree
    ic = dict(ic)
    ic["time_bin"] = np.zeros(len(ic["h"]), dtype=np.int8)
    be.build_from_cells(ic, cells, tasks, nt)
    be.integrator_init(**ip)
    log = []
    ti_current, ti_old = 0, 0
    for step in range(nsteps + 1):
        mab = get_max_active_bin(ti_current)
        be.set_time(ti_current, mab)
        dx = (0.0, 0.0)
        if step > 0:
            dx = be.drift(ti_old)
        be.step(do_sort=True)
        if step > 0:
            be.kick(2)
        ti_end_min, updated = be.timestep()
        lim = ()
        if limiter:
            # the wake-up flags come back only for the census of the log; the device keeps its own
            tb = be.state()["time_bin"]
            wake = be.runner_do_limiter_loop(want_wakeup=True)
            lim = limiter_census(wake, tb, mab)
            t_lim, nlim = be.runner_do_limiter()
            assert nlim == lim[0]
            ti_end_min = min(ti_end_min, t_lim)
        be.kick(1)
        log.append((ti_current, mab, updated, dx[0], dx[1], ti_end_min) + lim)
        if on_step is not None:
            on_step(step, _gpu_state(be, sphenix))
        ti_old, ti_current = ti_current, ti_end_min
    return _gpu_state(be, sphenix), log


def _gpu_state(be, sphenix):
    st = be.state()
    if sphenix:
        r = be.sphenix_results()
        st.update({"alpha_visc": r["alpha_visc"], "alpha_diff": r["alpha_diff"], "div_v_prev": r["div_v_previous_step"],
                   "div_v_dt": r["div_v_dt"], "laplace_u": r["laplace_u"]})
    return st
