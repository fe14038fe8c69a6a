"""Writes tests/golden/tree/*.npz: the cell tree and hydro task list the REFERENCE's own space_regrid,
space_split_recursive, engine_make_hydroloop_tasks_mapper and scheduler_splittasks build for seeded
particle sets (oracle/ref_tree.c; needs oracle/_ref, i.e. this container).  tests/test_tree.py holds the
oracle's restatement and the product's host builder to these fixtures."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import oracle_lib as ol  # noqa: E402
import tree_util as tu  # noqa: E402
from swift_subtask_dev_b200 import ics  # noqa: E402

CASES = {
    "uniform_split": (lambda: ics.uniform_random(40 ** 3, seed=5), dict(periodic=True, hscale=3.0)),
    "clustered_deep": (lambda: ics.clustered(30000, nclumps=16, seed=3), dict(periodic=True)),
    "open_box": (lambda: ics.uniform_random(28 ** 3, seed=9), dict(periodic=False)),
    "lattice": (lambda: ics.lattice(16), dict(periodic=True)),
}
for name, (gen, kw) in CASES.items():
    ic = gen()
    kw = dict(kw)
    if "hscale" in kw:   # larger h: fewer top-level cells, more particles in each, split (twice)
        ic["h"] = (ic["h"] * kw.pop("hscale")).astype(np.float32)
    prm = ol.make_params(**kw)
    order, cells, nc, cdim, tasks, nt = tu.reference_tree(prm, ic["x"], ic["h"])
    np.savez_compressed(os.path.join(HERE, "tree", name + ".npz"), x=ic["x"], h=ic["h"], params=repr(kw),
                        **tu.pack(cells, nc, tasks, nt, order, cdim))
    print(name, "cells", nc, "tasks", nt, "cdim", cdim)
