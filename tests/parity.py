"""Shared parity helpers: run the CUDA backend and the oracle on the same
input and compare (pair sets exactly, fields within the stated tolerance)."""
import numpy as np

import oracle_lib as ol
from swift_subtask_dev_b200 import HydroBackend, ics, make_params

# FP32 tolerances of BASELINE.json:north_star
TOL_RHO_H = 1e-5      # rel, rho and h
TOL_FORCE = 1e-4      # rel, accelerations, du/dt (and the other force-loop outputs)


def rel_err(a, b, floor):
    """max |a-b| / max(|b|, floor): relative error with the magnitude floor the
    reference's own tolerance files use (tests/difffloat.py row 3)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


def run_gpu(ic, prm_kwargs, record=False, pair_capacity=0):
    prm = make_params(**prm_kwargs)
    be = HydroBackend(prm)
    be.build(ic)
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


def run_oracle_on_tree(ic, tree, prm_kwargs, record=False, which="oracle", variant="minimal"):
    """Runs the oracle (or the compiled reference) on the GPU library's own tree
    and task list, so the comparison is on identical cells."""
    kw = dict(prm_kwargs)
    scheme = kw.pop("scheme", 0)
    kw.pop("til", None)
    kw.pop("idg_gamma", None)
    prm = ol.make_params(record=record, scheme=scheme, **kw)
    order = tree["order"]
    icc = ics.permute(ic, order)
    cells = (ol.Cell * tree["ncells"])()
    import ctypes as C
    C.memmove(cells, tree["cells"], C.sizeof(ol.Cell) * tree["ncells"])
    tasks = (ol.Task * max(tree["ntasks"], 1))()
    C.memmove(tasks, tree["tasks"], C.sizeof(ol.Task) * tree["ntasks"])
    o = ol.run_step(which, prm, icc, cells, tasks, tree["ntasks"], record=record, variant=variant)
    # back to the caller's particle order
    inv = np.empty_like(order)
    inv[order] = np.arange(len(order), dtype=order.dtype)
    for k in list(o.keys()):
        if isinstance(o[k], np.ndarray) and len(o[k]) == len(order):
            o[k] = o[k][inv]
    return o


def compare(gpu, orc, verbose=True):
    """Returns dict of error measures."""
    res = {}
    res["count_density"] = (gpu["counters"]["n_density"], orc["counts"][0])
    res["count_redo"] = (gpu["counters"]["n_density_redo"], orc["counts"][1])
    res["count_force"] = (gpu["counters"]["n_force"], orc["counts"][2])
    # first pass raw sums
    for k in ("d_rho", "d_rho_dh", "d_wcount", "d_wcount_dh", "d_div_v", "d_rot_v"):
        scale = float(np.abs(orc[k]).max()) or 1.0
        res[k] = rel_err(gpu[k], orc[k], 1e-3 * scale)
    res["h"] = rel_err(gpu["h"], orc["g_h"], 1e-30)
    res["rho"] = rel_err(gpu["rho"], orc["g_rho"], 1e-30)
    for kg, ko in (("pressure", "g_pressure"), ("soundspeed", "g_soundspeed"), ("f", "g_f"), ("balsara", "g_balsara")):
        scale = float(np.abs(orc[ko]).max()) or 1.0
        res[kg] = rel_err(gpu[kg], orc[ko], 1e-3 * scale)
    for kg, ko in (("a", "f_a"), ("u_dt", "f_u_dt"), ("h_dt", "f_h_dt"), ("v_sig", "f_v_sig")):
        # magnitude floor: 1% of the rms of the field (accelerations cancel to ~0 in uniform regions)
        rms = float(np.sqrt(np.mean(np.asarray(orc[ko], dtype=np.float64) ** 2))) or 1.0
        res[kg] = rel_err(gpu[kg], orc[ko], 1e-2 * rms)
    res["min_bin_mismatch"] = int(np.sum(gpu["min_ngb_time_bin"] != orc["f_min_ngb_time_bin"]))
    if verbose:
        for k, v in res.items():
            print("  %-16s %s" % (k, v))
    return res


def pair_set_diff(a, b):
    """Number of ordered pairs in exactly one of the two [m,2] lists."""
    sa = ol.sorted_pairs(np.asarray(a))
    sb = ol.sorted_pairs(np.asarray(b))
    if sa.shape == sb.shape and (sa == sb).all():
        return 0
    va = sa.view([("i", np.int64), ("j", np.int64)]).ravel()
    vb = sb.view([("i", np.int64), ("j", np.int64)]).ravel()
    return len(np.setxor1d(va, vb))
