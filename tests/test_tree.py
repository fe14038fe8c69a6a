"""The cell tree and the hydro task list, pinned against the REFERENCE ITSELF:
tests/golden/tree/*.npz hold what the reference's own space_regrid, space_split_recursive,
engine_make_hydroloop_tasks_mapper and scheduler_splittasks (scheduler_splittask_hydro) built for seeded
particle sets (oracle/ref_tree.c, tests/golden/make_golden_tree.py).  Both the oracle's restatement
(oracle_build_tree / oracle_build_tasks) and the product's host builder (csrc/host_tree.cpp, through
swift_cuda_host_plan: no GPU needed) must reproduce them: same cells (position, width, particle
range, split flag, parent, progeny), same particles in every leaf, same tasks."""
import ast
import glob
import os

import numpy as np
import pytest

import oracle_lib as ol
import tree_util as tu
from swift_subtask_dev_b200 import backend as be
from swift_subtask_dev_b200 import make_params

TREE_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tree")
NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(TREE_DIR, "*.npz")))


def _load(name):
    z = np.load(os.path.join(TREE_DIR, name + ".npz"))
    kw = ast.literal_eval(str(z["params"]))
    kw.pop("hscale", None)
    cells, nc, tasks, nt = tu.unpack(z)
    return z, kw, cells, nc, tasks, nt


@pytest.mark.parametrize("name", NAMES)
def test_oracle_tree_and_tasks_match_the_reference(name):
    z, kw, rcells, rnc, rtasks, rnt = _load(name)
    prm = ol.make_params(**kw)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, z["x"], z["h"])
    assert tuple(cdim) == tuple(int(v) for v in z["cdim"])
    got = tu.canonical(cells, len(cells), tasks, nt)
    ref = tu.canonical(rcells, rnc, rtasks, rnt)
    assert got[0] == ref[0], "cell tables differ"
    assert got[1] == ref[1], "task lists differ"
    a, b = tu.leaf_sets(cells, len(cells), order), tu.leaf_sets(rcells, rnc, z["order"])
    assert a.keys() == b.keys() and all(np.array_equal(a[k], b[k]) for k in a)


@pytest.mark.parametrize("name", NAMES)
def test_host_builder_tree_and_tasks_match_the_reference(name):
    z, kw, rcells, rnc, rtasks, rnt = _load(name)
    ic = {"x": z["x"], "h": z["h"]}
    plan = be.host_plan(make_params(**kw), ic)
    assert tuple(plan["cdim"]) == tuple(int(v) for v in z["cdim"])
    got = tu.canonical(plan["cells"], plan["ncells"], plan["tasks"], plan["ntasks"])
    ref = tu.canonical(rcells, rnc, rtasks, rnt)
    assert got[0] == ref[0], "cell tables differ"
    assert got[1] == ref[1], "task lists differ"
    a = tu.leaf_sets(plan["cells"], plan["ncells"], plan["order"])
    b = tu.leaf_sets(rcells, rnc, z["order"])
    assert a.keys() == b.keys() and all(np.array_equal(a[k], b[k]) for k in a)


@pytest.mark.skipif(not ol.ref_available("minimal"), reason="oracle/_ref not built (needs the reference tree)")
def test_live_reference_tree_on_a_fresh_particle_set():
    """Not only the fixtures: a particle set no fixture holds, built by the reference here and now."""
    from swift_subtask_dev_b200 import ics
    ic = ics.clustered(20000, nclumps=5, seed=77)
    kw = dict(periodic=True)
    prm = ol.make_params(**kw)
    rorder, rcells, rnc, rcdim, rtasks, rnt = tu.reference_tree(prm, ic["x"], ic["h"])
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    plan = be.host_plan(make_params(**kw), ic)
    ref = tu.canonical(rcells, rnc, rtasks, rnt)
    assert tu.canonical(cells, len(cells), tasks, nt) == ref
    assert tu.canonical(plan["cells"], plan["ncells"], plan["tasks"], plan["ntasks"]) == ref
