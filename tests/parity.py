"""Shared parity helpers: run the CUDA backend and the oracle on the same
input and compare.  The CUDA path sums every particle's neighbours in the
order the reference's single-threaded runner does and uses the same IEEE
operations (no FMA contraction), so the bar is BIT-EXACT: pair lists, counts
and every per-particle field.  The tolerances of BASELINE.json:north_star are
kept as named constants for the one place they are still needed (Tillotson
pressure / sound speed go through expf, which differs in the last ulp between
glibc and CUDA)."""
import ctypes as C

import numpy as np

import oracle_lib as ol
from swift_subtask_dev_b200 import HydroBackend, ics, make_params
from swift_subtask_dev_b200 import backend as be_mod

# FP32 tolerances of BASELINE.json:north_star
TOL_RHO_H = 1e-5      # rel, rho and h
TOL_FORCE = 1e-4      # rel, accelerations, du/dt (and the other force-loop outputs)

# (GPU result key, oracle key)
FIELD_MAP = [("d_rho", "d_rho"), ("d_rho_dh", "d_rho_dh"), ("d_wcount", "d_wcount"), ("d_wcount_dh", "d_wcount_dh"),
             ("d_div_v", "d_div_v"), ("d_rot_v", "d_rot_v"),
             ("h", "g_h"), ("rho", "g_rho"), ("f", "g_f"), ("pressure", "g_pressure"), ("soundspeed", "g_soundspeed"),
             ("balsara", "g_balsara"),
             ("a", "f_a"), ("u_dt", "f_u_dt"), ("h_dt", "f_h_dt"), ("v_sig", "f_v_sig"),
             ("min_ngb_time_bin", "f_min_ngb_time_bin")]


def rel_err(a, b, floor):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


def gpu_params(kw, eos=None):
    """make_params of the backend from the keyword set the oracle uses (+ Tillotson table)."""
    kw = dict(kw)
    if eos is not None:
        til = []
        for k in range(eos.n_til):
            t = be_mod.Til()
            for name, _ in be_mod.Til._fields_:
                setattr(t, name, getattr(eos.til[k], name))
            til.append(t)
        kw["til"] = til
        kw["idg_gamma"] = tuple(eos.idg_gamma)
    return make_params(**kw)


def run_gpu(ic, prm_kwargs, record=False, pair_capacity=0, tree=None, eos=None, fields_mode=0):
    """tree=None: the library bins the particles itself (build_from_particles);
    tree=(cells, tasks, ntasks): the drop-in path, caller-built tree (build_from_cells)."""
    prm = gpu_params(prm_kwargs, eos)
    prm.fields_mode = fields_mode
    be = HydroBackend(prm)
    if tree is None:
        be.build(ic)
    else:
        cells, tasks, nt = tree
        gc = (be_mod.Cell * len(cells))()
        C.memmove(gc, cells, C.sizeof(be_mod.Cell) * len(cells))
        gt = (be_mod.Task * max(nt, 1))()
        C.memmove(gt, tasks, C.sizeof(be_mod.Task) * nt)
        be.build_from_cells(ic, gc, gt, nt)
    if record:
        be.record_pairs(pair_capacity)
    be.step()
    out = be.results(first_pass=True)
    out["counters"] = be.counters()
    out["tree"] = be.tree()
    if record:
        out["pairs"] = [be.fetch_pairs(k, pair_capacity) for k in range(3)]
    be.close()
    return out


def run_oracle_on_tree(ic, tree, prm_kwargs, record=False, which="oracle", variant="minimal", eos=None):
    """Runs the oracle (or the compiled reference) on the GPU library's own tree
    and task list, so the comparison is on identical cells."""
    kw = dict(prm_kwargs)
    prm = ol.make_params(record=record, **kw)
    order = tree["order"]
    icc = ics.permute(ic, order)
    cells = (ol.Cell * tree["ncells"])()
    C.memmove(cells, tree["cells"], C.sizeof(ol.Cell) * tree["ncells"])
    tasks = (ol.Task * max(tree["ntasks"], 1))()
    C.memmove(tasks, tree["tasks"], C.sizeof(ol.Task) * tree["ntasks"])
    o = ol.run_step(which, prm, icc, cells, tasks, tree["ntasks"], record=record, variant=variant, eos=eos)
    inv = np.empty_like(order)
    inv[order] = np.arange(len(order), dtype=order.dtype)
    for k in list(o.keys()):
        if isinstance(o[k], np.ndarray) and len(o[k]) == len(order):
            o[k] = o[k][inv]
    return o


def compare(gpu, orc, active=None, verbose=True):
    """Returns {field: (n_mismatching_particles, max_abs_diff)} plus the counts."""
    res = {}
    res["count_density"] = (gpu["counters"]["n_density"], orc["counts"][0])
    res["count_redo"] = (gpu["counters"]["n_density_redo"], orc["counts"][1])
    res["count_force"] = (gpu["counters"]["n_force"], orc["counts"][2])
    for kg, ko in FIELD_MAP:
        a, b = np.asarray(gpu[kg]), np.asarray(orc[ko])
        if active is not None:
            a, b = a[active], b[active]
        bad = a != b
        if a.dtype.kind == "f":
            bad &= ~(np.isnan(a) & np.isnan(b))
        nbad = int(np.count_nonzero(bad.reshape(len(a), -1).any(axis=1))) if a.size else 0
        diff = float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64)))) if nbad else 0.0
        res[kg] = (nbad, diff)
    if verbose:
        for k, v in res.items():
            print("  %-18s %s" % (k, v))
    return res


# what the FAST fields mode (swift_cuda_params.fields_mode = 1) may change: the force loop's outputs
FAST_FIELDS = ("a", "u_dt", "h_dt", "v_sig")


def assert_fast(gpu, orc, res, active=None):
    """FAST mode against the reference: counts, the neighbour sets (checked by the caller) and every
    field up to and including the ghost's are bit-exact; the force loop's fields agree within
    TOL_FORCE = 1e-4 (BASELINE.json: north_star), relative to max(|reference|, 0.1 x the rms of the field
    over the particles): a particle whose acceleration cancels to ~0 has no meaningful relative error."""
    assert_exact(res, skip=FAST_FIELDS)
    worst = {}
    for kg, ko in FIELD_MAP:
        if kg not in FAST_FIELDS:
            continue
        a = np.asarray(gpu[kg], dtype=np.float64)
        b = np.asarray(orc[ko], dtype=np.float64)
        if active is not None:
            a, b = a[active], b[active]
        rms = float(np.sqrt(np.mean(b ** 2))) if b.size else 0.0
        err = np.abs(a - b) / np.maximum(np.abs(b), max(0.1 * rms, 1e-30))
        worst[kg] = float(err.max()) if err.size else 0.0
        assert worst[kg] <= TOL_FORCE, "FAST field %s: relative error %g > %g" % (kg, worst[kg], TOL_FORCE)
    return worst


def assert_exact(res, skip=()):
    for k in ("count_density", "count_redo", "count_force"):
        assert res[k][0] == res[k][1], (k, res[k])
    for kg, _ in FIELD_MAP:
        if kg in skip:
            continue
        assert res[kg][0] == 0, "field %s differs on %d particles (max abs %g)" % (kg, res[kg][0], res[kg][1])


def pair_set_diff(a, b):
    """Number of ordered pairs in exactly one of the two [m,2] lists."""
    sa = ol.sorted_pairs(np.asarray(a))
    sb = ol.sorted_pairs(np.asarray(b))
    if sa.shape == sb.shape and (sa == sb).all():
        return 0
    va = np.ascontiguousarray(sa).view([("i", np.int64), ("j", np.int64)]).ravel()
    vb = np.ascontiguousarray(sb).view([("i", np.int64), ("j", np.int64)]).ravel()
    return len(np.setxor1d(va, vb))
