"""Loading of the committed golden fixtures (tests/golden/*.npz, written by
tests/golden/make_golden.py from the reference's own compiled operators)."""
import ast
import glob
import hashlib
import os

import numpy as np

import oracle_lib as ol

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["params"] = ast.literal_eval(str(g["params"]))
    g["variant"] = str(g["variant"])
    g["ic"] = {"x": g["x"], "v": g["v"], "mass": g["mass"], "h": g["h"], "u": g["u"], "id": g["id"],
               "mat_id": g.get("mat_id"), "time_bin": g.get("time_bin")}
    for k in ("in_rho", "in_pressure", "in_f", "in_soundspeed", "in_balsara"):
        g["ic"][k] = g.get(k)
    return g


SPHENIX_DIR = os.path.join(GOLDEN_DIR, "sphenix")


def sphenix_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(SPHENIX_DIR, "*.npz")))


def load_sphenix(name):
    """A fixture of the three-loop scheme (tests/golden/make_golden_sphenix.py): as load(), plus the hydro_props
    of the scheme (g["props"]) and the viscosity / diffusion state among the inputs."""
    z = np.load(os.path.join(SPHENIX_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["params"] = ast.literal_eval(str(g["params"]))
    g["props"] = ast.literal_eval(str(g["props"]))
    g["variant"] = str(g["variant"])
    g["ic"] = {"x": g["x"], "v": g["v"], "mass": g["mass"], "h": g["h"], "u": g["u"], "id": g["id"],
               "mat_id": None, "time_bin": g.get("time_bin")}
    for k in ("in_rho", "in_pressure", "in_f", "in_soundspeed", "in_balsara") + tuple(ol.SPHENIX_IN):
        g["ic"][k] = g.get(k)
    return g


def tree_structs(g, cell_cls=ol.Cell, task_cls=ol.Task):
    """ctypes arrays of the stored tree and task list."""
    n = len(g["cell_ints"])
    cells = (cell_cls * n)()
    for k in range(n):
        c = cells[k]
        c.loc[:] = g["cell_loc"][k]
        c.width[:] = g["cell_width"][k]
        first, count, parent, depth, split = (int(v) for v in g["cell_ints"][k][:5])
        c.first, c.count, c.parent, c.depth, c.split = first, count, parent, depth, split
        c.progeny[:] = [int(v) for v in g["cell_ints"][k][5:13]]
    nt = len(g["tasks"])
    tasks = (task_cls * max(nt, 1))()
    for k in range(nt):
        tasks[k].type, tasks[k].ci, tasks[k].cj = (int(v) for v in g["tasks"][k])
    return cells, tasks, nt


def pair_digest(pairs):
    p = ol.sorted_pairs(np.asarray(pairs, dtype=np.int64))
    return hashlib.sha256(np.ascontiguousarray(p).tobytes()).hexdigest()


def ngb_counts(pairs, ids):
    order = np.argsort(ids)
    out = np.zeros(len(ids), dtype=np.int32)
    if len(pairs):
        u, cnt = np.unique(pairs[:, 0], return_counts=True)
        pos = order[np.searchsorted(ids[order], u)]
        out[pos] = cnt
    return out
