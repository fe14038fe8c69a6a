"""ctypes bindings of the TEST-ONLY checkers: oracle/liboracle.so (plain-C
restatement) and oracle/_ref/libswiftref_*.so (the reference's own compiled
operators behind oracle/ref_harness.c).  Never imported by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")


class Cell(C.Structure):
    _fields_ = [("loc", C.c_double * 3), ("width", C.c_double * 3), ("first", C.c_int), ("count", C.c_int),
                ("parent", C.c_int), ("progeny", C.c_int * 8), ("depth", C.c_int), ("split", C.c_int)]


class Task(C.Structure):
    _fields_ = [("type", C.c_int), ("ci", C.c_int), ("cj", C.c_int)]


class Params(C.Structure):
    _fields_ = [("dim", C.c_double * 3), ("periodic", C.c_int), ("eta_neighbours", C.c_float),
                ("h_tolerance", C.c_float), ("h_max", C.c_float), ("h_min", C.c_float),
                ("max_smoothing_iterations", C.c_int), ("viscosity_alpha", C.c_float),
                ("max_active_bin", C.c_int), ("record_pairs", C.c_int), ("scheme", C.c_int)]


class Til(C.Structure):
    _fields_ = [(k, C.c_float) for k in
                ("rho_0", "a", "b", "A", "B", "u_0", "u_iv", "u_cv", "alpha", "beta", "eta_min", "eta_zero", "P_min")] \
        + [("mat_id", C.c_int)]


class EosTable(C.Structure):
    # oracle_eos_table (oracle/oracle_api.h) == swift_cuda_eos_table, field for field
    _fields_ = [("kind", C.c_int), ("mat_id", C.c_int), ("num_rho", C.c_int), ("num_T", C.c_int),
                ("log_rho", C.POINTER(C.c_float)), ("log_u", C.POINTER(C.c_float)), ("P", C.POINTER(C.c_float)),
                ("c", C.POINTER(C.c_float)), ("P_tiny", C.c_float), ("c_tiny", C.c_float),
                ("log_rho_min", C.c_float), ("log_rho_step", C.c_float), ("inv_log_rho_step", C.c_float),
                ("log_u_min", C.c_float), ("log_u_step", C.c_float), ("inv_log_u_step", C.c_float),
                ("bulk_mod", C.c_float), ("P_min_for_c_min", C.c_float)]


class Eos(C.Structure):
    _fields_ = [("n_til", C.c_int), ("til", Til * 8), ("idg_gamma", C.c_float * 4),
                ("n_tab", C.c_int), ("tab", C.POINTER(EosTable))]


_FP = C.POINTER(C.c_float)


class Parts(C.Structure):
    _fields_ = [("n", C.c_int), ("x", C.POINTER(C.c_double)), ("v", _FP), ("mass", _FP), ("h", _FP), ("u", _FP),
                ("mat_id", C.POINTER(C.c_int)), ("id", C.POINTER(C.c_longlong)), ("time_bin", C.POINTER(C.c_byte)),
                ("d_rho", _FP), ("d_rho_dh", _FP), ("d_wcount", _FP), ("d_wcount_dh", _FP), ("d_div_v", _FP),
                ("d_rot_v", _FP),
                ("g_h", _FP), ("g_rho", _FP), ("g_f", _FP), ("g_pressure", _FP), ("g_soundspeed", _FP),
                ("g_balsara", _FP),
                ("f_a", _FP), ("f_u_dt", _FP), ("f_h_dt", _FP), ("f_v_sig", _FP),
                ("f_min_ngb_time_bin", C.POINTER(C.c_byte)),
                ("in_rho", _FP), ("in_pressure", _FP), ("in_f", _FP), ("in_soundspeed", _FP), ("in_balsara", _FP),
                ("in_h_build", _FP),
                ("l_time_bin", C.POINTER(C.c_byte)), ("l_wakeup", C.POINTER(C.c_byte)), ("l_only", C.c_int)]


class PairLog(C.Structure):
    _fields_ = [("pairs", C.POINTER(C.c_longlong) * 3), ("capacity", C.c_longlong * 3),
                ("count", C.c_longlong * 3), ("ghost_sweeps", C.c_int)]


class Sphenix(C.Structure):
    """oracle_sphenix (oracle/oracle_api.h): what the three-loop scheme adds to Params / Parts / PairLog."""
    _fields_ = [("viscosity_alpha_max", C.c_float), ("viscosity_alpha_min", C.c_float),
                ("viscosity_length", C.c_float), ("diffusion_beta", C.c_float), ("diffusion_alpha_max", C.c_float),
                ("diffusion_alpha_min", C.c_float), ("time_base", C.c_double),
                ("in_alpha_visc", _FP), ("in_alpha_diff", _FP), ("in_div_v_prev", _FP),
                ("g_div_v", _FP), ("r_v_sig", _FP), ("r_laplace_u", _FP), ("r_alpha_visc_max_ngb", _FP),
                ("e_laplace_u", _FP), ("e_alpha_visc", _FP), ("e_alpha_diff", _FP), ("e_div_v_prev", _FP),
                ("e_div_v_dt", _FP),
                ("grad_pairs", C.POINTER(C.c_longlong)), ("grad_capacity", C.c_longlong), ("grad_count", C.c_longlong)]


SPHENIX_OUT = ("g_div_v", "r_v_sig", "r_laplace_u", "r_alpha_visc_max_ngb", "e_laplace_u", "e_alpha_visc",
               "e_alpha_diff", "e_div_v_prev", "e_div_v_dt")
SPHENIX_IN = ("in_alpha_visc", "in_alpha_diff", "in_div_v_prev")
# hydro_props defaults of the scheme (src/hydro/SPHENIX/hydro_parameters.h:53-101) and the harness's time base
SPHENIX_PROPS = dict(viscosity_alpha=0.1, viscosity_alpha_max=2.0, viscosity_alpha_min=0.0, viscosity_length=0.05,
                     diffusion_beta=1.0, diffusion_alpha_max=1.0, diffusion_alpha_min=0.0, time_base=1e-3)

OUT_FIELDS = {"d_rho": 1, "d_rho_dh": 1, "d_wcount": 1, "d_wcount_dh": 1, "d_div_v": 1, "d_rot_v": 3,
              "g_h": 1, "g_rho": 1, "g_f": 1, "g_pressure": 1, "g_soundspeed": 1, "g_balsara": 1,
              "f_a": 3, "f_u_dt": 1, "f_h_dt": 1, "f_v_sig": 1}

_libs = {}


def build_oracle():
    subprocess.run(["make", "-C", ORACLE_DIR, "liboracle.so"], check=True, capture_output=True)


def oracle():
    if "oracle" not in _libs:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        lib = C.CDLL(path)
        lib.oracle_build_tree.restype = C.c_int
        lib.oracle_build_tasks.restype = C.c_int
        lib.oracle_run_step.restype = C.c_int
        lib.oracle_run_step_sphenix.restype = C.c_int
        lib.oracle_free.argtypes = [C.c_void_p]
        _libs["oracle"] = lib
    return _libs["oracle"]


def ref_available(variant="minimal"):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libswiftref_%s.so" % variant))


def ref(variant="minimal"):
    key = "ref_" + variant
    if key not in _libs:
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libswiftref_%s.so" % variant))
        lib.ref_run_step.restype = C.c_int
        lib.ref_run_step_sphenix.restype = C.c_int
        lib.ref_get_eos.restype = C.c_int
        _libs[key] = lib
    return _libs[key]


def ref_eos():
    """The reference's own Tillotson / ideal-gas parameters in internal units
    (needs the compiled planetary reference)."""
    e = Eos()
    rc = ref("planetary").ref_get_eos(C.byref(e))
    if rc != 0:
        raise RuntimeError("ref_get_eos failed")
    return e


def eos_from_json(path=os.path.join(ROOT, "tests", "golden", "eos_planetary.json")):
    """The same parameters from the committed fixture (GPU box: no reference there)."""
    import json
    with open(path) as f:
        d = json.load(f)
    e = Eos()
    e.n_til = len(d["til"])
    for k, t in enumerate(d["til"]):
        for name, _ in Til._fields_:
            setattr(e.til[k], name, t[name])
    e.idg_gamma[:] = d["idg_gamma"]
    return e


def eos_to_dict(e):
    return {"til": [{name: getattr(e.til[k], name) for name, _ in Til._fields_} for k in range(e.n_til)],
            "idg_gamma": list(e.idg_gamma)}


def make_params(dim=(1.0, 1.0, 1.0), periodic=True, eta=1.2348, h_tolerance=1e-4, h_max=1e10, h_min=0.0,
                max_iter=30, alpha=0.8, max_active_bin=56, record=False, scheme=0):
    p = Params()
    p.dim[:] = dim
    p.periodic = int(periodic)
    p.eta_neighbours = eta
    p.h_tolerance = h_tolerance
    p.h_max = h_max
    p.h_min = h_min
    p.max_smoothing_iterations = max_iter
    p.viscosity_alpha = alpha
    p.max_active_bin = max_active_bin
    p.record_pairs = int(record)
    p.scheme = scheme
    return p


def build_tree_and_tasks(prm, x, h):
    """Returns (order, cells ndarray-of-struct ctypes array, cdim, tasks ctypes array)."""
    lib = oracle()
    n = len(h)
    x = np.ascontiguousarray(x, dtype=np.float64)
    h = np.ascontiguousarray(h, dtype=np.float32)
    order = np.empty(n, dtype=np.int32)
    cells_p = C.POINTER(Cell)()
    ncells = C.c_int()
    cdim = (C.c_int * 3)()
    rc = lib.oracle_build_tree(C.byref(prm), n, x.ctypes.data_as(C.c_void_p), h.ctypes.data_as(C.c_void_p),
                               order.ctypes.data_as(C.c_void_p), C.byref(cells_p), C.byref(ncells), cdim)
    if rc != 0:
        raise RuntimeError("oracle_build_tree failed rc=%d" % rc)
    cells = (Cell * ncells.value)()
    C.memmove(cells, cells_p, C.sizeof(Cell) * ncells.value)
    lib.oracle_free(cells_p)
    hs = np.ascontiguousarray(h[order])
    tasks_p = C.POINTER(Task)()
    ntasks = C.c_int()
    lib.oracle_build_tasks(C.byref(prm), ncells.value, cells, cdim, hs.ctypes.data_as(C.c_void_p),
                           C.byref(tasks_p), C.byref(ntasks))
    tasks = (Task * max(ntasks.value, 1))()
    C.memmove(tasks, tasks_p, C.sizeof(Task) * ntasks.value)
    lib.oracle_free(tasks_p)
    return order, cells, tuple(cdim), tasks, ntasks.value


def run_step(which, prm, ic, cells, tasks, ntasks, record=False, pair_capacity=None, eos=None, variant="minimal",
             sphenix=None):
    """ic: dict with x[n,3] f64, v[n,3], mass, h, u (f32), id (i64), optional mat_id, time_bin -- already in CELL ORDER.
    which: 'oracle' or 'ref'.  Returns dict of outputs (+ 'pairs', 'counts', 'ghost_sweeps').
    ic["l_time_bin"] (int8, the bins after the time-step task): the time-step limiter loop runs after end_force and
    out["l_wakeup"] is p->limiter_data.wakeup of every particle; with ic["l_only"] nothing but sort + that loop runs.
    sphenix: dict of the SPHENIX hydro_props (SPHENIX_PROPS keys; prm.scheme must be 2): runs the three-loop
    scheme; ic may hold in_alpha_visc / in_alpha_diff / in_div_v_prev; the outputs gain SPHENIX_OUT,
    counts[3] and pairs[3] (the gradient loop)."""
    n = len(ic["h"])
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dtype=dt)
        keep.append(a)
        return a

    P = Parts()
    P.n = n
    x = arr(ic["x"], np.float64)
    P.x = x.ctypes.data_as(C.POINTER(C.c_double))
    for k in ("v", "mass", "h", "u"):
        a = arr(ic[k], np.float32)
        setattr(P, k, a.ctypes.data_as(_FP))
    ids = arr(ic["id"], np.int64)
    P.id = ids.ctypes.data_as(C.POINTER(C.c_longlong))
    if ic.get("mat_id") is not None:
        m = arr(ic["mat_id"], np.int32)
        P.mat_id = m.ctypes.data_as(C.POINTER(C.c_int))
    if ic.get("time_bin") is not None:
        tb = arr(ic["time_bin"], np.int8)
        P.time_bin = tb.ctypes.data_as(C.POINTER(C.c_byte))
    for k in ("in_rho", "in_pressure", "in_f", "in_soundspeed", "in_balsara", "in_h_build"):
        if ic.get(k) is not None:
            setattr(P, k, arr(ic[k], np.float32).ctypes.data_as(_FP))
    out = {}
    if ic.get("l_time_bin") is not None:
        P.l_time_bin = arr(ic["l_time_bin"], np.int8).ctypes.data_as(C.POINTER(C.c_byte))
        out["l_wakeup"] = np.zeros(n, dtype=np.int8)
        P.l_wakeup = out["l_wakeup"].ctypes.data_as(C.POINTER(C.c_byte))
        P.l_only = int(bool(ic.get("l_only")))
    for k, w in OUT_FIELDS.items():
        out[k] = np.zeros((n, w) if w > 1 else n, dtype=np.float32)
        setattr(P, k, out[k].ctypes.data_as(_FP))
    out["f_min_ngb_time_bin"] = np.zeros(n, dtype=np.int8)
    P.f_min_ngb_time_bin = out["f_min_ngb_time_bin"].ctypes.data_as(C.POINTER(C.c_byte))
    log = PairLog()
    bufs = []
    prm.record_pairs = int(record)
    if record:
        cap = pair_capacity or (n * 400)
        for s in range(3):
            b = np.zeros((cap, 2), dtype=np.int64)
            bufs.append(b)
            log.pairs[s] = b.ctypes.data_as(C.POINTER(C.c_longlong))
            log.capacity[s] = cap
    ncells = len(cells)
    if sphenix is not None:
        S = Sphenix()
        for k in ("viscosity_alpha_max", "viscosity_alpha_min", "viscosity_length", "diffusion_beta",
                  "diffusion_alpha_max", "diffusion_alpha_min", "time_base"):
            setattr(S, k, sphenix[k])
        for k in SPHENIX_IN:
            if ic.get(k) is not None:
                setattr(S, k, arr(ic[k], np.float32).ctypes.data_as(_FP))
        for k in SPHENIX_OUT:
            out[k] = np.zeros(n, dtype=np.float32)
            setattr(S, k, out[k].ctypes.data_as(_FP))
        gb = None
        if record:
            gb = np.zeros((pair_capacity or (n * 400), 2), dtype=np.int64)
            S.grad_pairs = gb.ctypes.data_as(C.POINTER(C.c_longlong))
            S.grad_capacity = len(gb)
        if which == "oracle":
            rc = oracle().oracle_run_step_sphenix(C.byref(prm), C.byref(P), ncells, cells, ntasks, tasks, C.byref(log),
                                                  C.byref(S))
        else:
            rc = ref("sphenix").ref_run_step_sphenix(C.byref(prm), C.byref(P), ncells, cells, ntasks, tasks,
                                                     C.byref(log), C.byref(S))
        if rc != 0:
            raise RuntimeError("%s run_step_sphenix rc=%d" % (which, rc))
        out["counts"] = [int(log.count[s]) for s in range(3)] + [int(S.grad_count)]
        out["ghost_sweeps"] = int(log.ghost_sweeps)
        if record:
            assert S.grad_count <= S.grad_capacity, "pair buffer too small"
            out["pairs"] = []
            for s in range(3):
                assert log.count[s] <= log.capacity[s], "pair buffer too small"
                out["pairs"].append(bufs[s][:log.count[s]].copy())
            out["pairs"].append(gb[:S.grad_count].copy())
        return out
    if which == "oracle":
        e = eos if eos is not None else Eos()
        rc = oracle().oracle_run_step(C.byref(prm), C.byref(e), C.byref(P), ncells, cells, ntasks, tasks, C.byref(log))
    else:
        rc = ref(variant).ref_run_step(C.byref(prm), C.byref(P), ncells, cells, ntasks, tasks, C.byref(log))
    if rc != 0:
        raise RuntimeError("%s run_step rc=%d" % (which, rc))
    out["counts"] = [int(log.count[s]) for s in range(3)]
    out["ghost_sweeps"] = int(log.ghost_sweeps)
    if record:
        out["pairs"] = []
        for s in range(3):
            assert log.count[s] <= log.capacity[s], "pair buffer too small"
            out["pairs"].append(bufs[s][:log.count[s]].copy())
    return out


def sorted_pairs(p):
    """Canonical (lexicographically sorted) copy of an [m,2] pair list."""
    if len(p) == 0:
        return p
    idx = np.lexsort((p[:, 1], p[:, 0]))
    return p[idx]
