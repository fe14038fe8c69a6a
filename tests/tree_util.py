"""Canonical form of a cell tree + hydro task list (cells named by depth and position, so that the
numbering of progeny and the order of tasks do not matter) and the call into the reference's own
tree / task code (oracle/ref_tree.c)."""
import ctypes as C

import numpy as np

import oracle_lib as ol


def canonical(cells, ncells, tasks, ntasks, order=None):
    """-> (cell table {name: (width, first, count, split, parent name)}, sorted task list)."""
    def name(k):
        c = cells[k]
        return (int(c.depth),) + tuple(float(v) for v in c.loc)
    table = {}
    for k in range(ncells):
        c = cells[k]
        if c.count == 0:
            continue
        table[name(k)] = (tuple(float(v) for v in c.width), int(c.first), int(c.count), int(bool(c.split)),
                          name(c.parent) if c.parent >= 0 else None,
                          tuple(sorted(name(p) for p in c.progeny if p >= 0 and cells[p].count > 0)))
    tl = []
    for k in range(ntasks):
        t = tasks[k]
        a, b = name(t.ci), (name(t.cj) if t.cj >= 0 else None)
        if b is not None and b < a:
            a, b = b, a
        tl.append((int(t.type), a, b))
    return table, sorted(tl, key=repr)


def leaf_sets(cells, ncells, order):
    """{leaf name: sorted input indices of its particles}"""
    out = {}
    for k in range(ncells):
        c = cells[k]
        if c.split or c.count == 0:
            continue
        out[(int(c.depth),) + tuple(float(v) for v in c.loc)] = np.sort(order[c.first:c.first + c.count])
    return out


def reference_tree(prm, x, h):
    """The reference's own space_regrid + space_split_recursive + engine_make_hydroloop_tasks_mapper +
    scheduler_splittasks on these particles (needs oracle/_ref)."""
    lib = ol.ref("minimal")
    n = len(h)
    x = np.ascontiguousarray(x, dtype=np.float64)
    h = np.ascontiguousarray(h, dtype=np.float32)
    order = np.empty(n, dtype=np.int32)
    cells_p, tasks_p = C.POINTER(ol.Cell)(), C.POINTER(ol.Task)()
    nc, nt = C.c_int(), C.c_int()
    cdim = (C.c_int * 3)()
    rc = lib.ref_build_tree_tasks(C.byref(prm), n, x.ctypes.data_as(C.c_void_p), h.ctypes.data_as(C.c_void_p),
                                  order.ctypes.data_as(C.c_void_p), C.byref(cells_p), C.byref(nc), cdim,
                                  C.byref(tasks_p), C.byref(nt))
    if rc != 0:
        raise RuntimeError("ref_build_tree_tasks failed rc=%d" % rc)
    cells = (ol.Cell * nc.value)()
    C.memmove(cells, cells_p, C.sizeof(ol.Cell) * nc.value)
    tasks = (ol.Task * max(nt.value, 1))()
    C.memmove(tasks, tasks_p, C.sizeof(ol.Task) * nt.value)
    return order, cells, nc.value, tuple(cdim), tasks, nt.value


def pack(cells, ncells, tasks, ntasks, order, cdim):
    """numpy arrays for an .npz fixture."""
    ci = np.zeros((ncells, 13), dtype=np.int64)
    loc = np.zeros((ncells, 3))
    wid = np.zeros((ncells, 3))
    for k in range(ncells):
        c = cells[k]
        loc[k], wid[k] = list(c.loc), list(c.width)
        ci[k, :5] = (c.first, c.count, c.parent, c.depth, c.split)
        ci[k, 5:] = list(c.progeny)
    tk = np.array([(tasks[k].type, tasks[k].ci, tasks[k].cj) for k in range(ntasks)], dtype=np.int64).reshape(-1, 3)
    return dict(cell_loc=loc, cell_width=wid, cell_ints=ci, tasks=tk, order=np.asarray(order), cdim=np.array(cdim))


def unpack(z):
    n = len(z["cell_ints"])
    cells = (ol.Cell * n)()
    for k in range(n):
        c = cells[k]
        c.loc[:] = z["cell_loc"][k]
        c.width[:] = z["cell_width"][k]
        c.first, c.count, c.parent, c.depth, c.split = (int(v) for v in z["cell_ints"][k][:5])
        c.progeny[:] = [int(v) for v in z["cell_ints"][k][5:13]]
    nt = len(z["tasks"])
    tasks = (ol.Task * max(nt, 1))()
    for k in range(nt):
        tasks[k].type, tasks[k].ci, tasks[k].cj = (int(v) for v in z["tasks"][k])
    return cells, n, tasks, nt
