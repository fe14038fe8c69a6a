"""CPU tests of the checker itself: the plain-C oracle against the golden
vectors produced by the reference's own compiled operators, and (when
oracle/_ref is present) against the reference live on more inputs."""
import numpy as np
import pytest

import golden_util as gu
import oracle_lib as ol
from swift_subtask_dev_b200 import ics

FIELDS = list(ol.OUT_FIELDS) + ["f_min_ngb_time_bin"]


def _active(g):
    """Outputs are defined for ACTIVE particles only (the others are not touched by the step)."""
    if g.get("time_bin") is None:
        return np.ones(len(g["h"]), dtype=bool)
    return g["time_bin"] <= g["params"].get("max_active_bin", 56)


def _eos(variant):
    return ol.eos_from_json() if variant == "planetary" else None


@pytest.mark.parametrize("name", gu.names())
def test_oracle_reproduces_reference_golden(name):
    g = gu.load(name)
    cells, tasks, nt = gu.tree_structs(g)
    prm = ol.make_params(record=True, **g["params"])
    o = ol.run_step("oracle", prm, g["ic"], cells, tasks, nt, record=True, eos=_eos(g["variant"]))
    assert o["counts"] == [int(v) for v in g["counts"]]
    for s in range(3):
        assert gu.pair_digest(o["pairs"][s]) == str(g["digests"][s]), "pair list of loop %d differs" % s
        assert np.array_equal(gu.ngb_counts(o["pairs"][s], g["id"]), g["ngb%d" % s])
    act = _active(g)
    for k in FIELDS:
        assert np.array_equal(o[k][act], g[k][act]), "field %s is not bit-identical to the reference's" % k


@pytest.mark.parametrize("name", gu.names())
def test_oracle_tree_matches_golden(name):
    """The tree and the task list stored with the golden vectors are the ones
    the oracle's builder (space_regrid / space_split / scheduler_splittask) makes."""
    g = gu.load(name)
    prm = ol.make_params(**g["params"])
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, g["x"], g["h"])
    assert np.array_equal(order, np.arange(len(order))), "inputs are stored in cell order"
    assert tuple(cdim) == tuple(int(v) for v in g["cdim"])
    assert len(cells) == len(g["cell_ints"]) and nt == len(g["tasks"])
    got = np.array([[c.first, c.count, c.parent, c.depth, c.split] + list(c.progeny) for c in cells], dtype=np.int32)
    assert np.array_equal(got, g["cell_ints"])
    assert np.array_equal(np.array([list(c.loc) for c in cells]), g["cell_loc"])
    assert np.array_equal(np.array([[tasks[k].type, tasks[k].ci, tasks[k].cj] for k in range(nt)], dtype=np.int32),
                          g["tasks"])


LIVE = [
    ("uniform20", lambda: ics.uniform_random(20 ** 3, seed=2), dict(periodic=True), "minimal"),
    ("uniform_split", lambda: ics.uniform_random(40000, seed=4), dict(periodic=True), "minimal"),
    ("clustered", lambda: ics.clustered(20000, nclumps=8, seed=6), dict(periodic=False), "minimal"),
    ("lattice_nojitter", lambda: ics.lattice(12, jitter=0.0), dict(periodic=True), "minimal"),
]


@pytest.mark.parametrize("name,make,kw,variant", LIVE, ids=[c[0] for c in LIVE])
def test_oracle_vs_compiled_reference(name, make, kw, variant):
    if not ol.ref_available(variant):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    ic = make()
    prm = ol.make_params(record=True, **kw)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    icc = ics.permute(ic, order)
    o = ol.run_step("oracle", prm, icc, cells, tasks, nt, record=True)
    r = ol.run_step("ref", prm, icc, cells, tasks, nt, record=True, variant=variant)
    assert o["counts"] == r["counts"]
    for s in range(3):
        # the SET of ordered pairs is always identical; their order too unless two
        # particles of a cell share a sort key (tie order is unspecified in the reference)
        assert gu.pair_digest(o["pairs"][s]) == gu.pair_digest(r["pairs"][s])
    exact = all(np.array_equal(o["pairs"][s], r["pairs"][s]) for s in range(3))
    for k in FIELDS:
        if exact:
            assert np.array_equal(o[k], r[k]), k
        else:
            a, b = o[k].astype(np.float64), r[k].astype(np.float64)
            # the grad-h term f is a cancelling sum with the dimension of a mass
            scale = float(icc["mass"].max()) if k == "g_f" else (np.abs(b).max() or 1.0)
            assert np.abs(a - b).max() <= 2e-5 * scale, k


def test_kernel_known_answers():
    """Known answers the reference prints/asserts: neighbour target 48.0009 for
    eta = 1.23485 (tests/test27cells.c:460), kernel_root and kernel_norm
    (src/kernel_hydro.h:205-241), W >= 0 and dW <= 0 (tests/testKernel.c)."""
    kg = np.float32(1.825742)
    kroot = np.float32(0.5) * np.float32(16.0 / np.pi) * np.float32(1.0 / (float(kg) ** 3))
    assert abs(float(kroot) - 0.418429106) < 1e-7
    knorm = 4.0 * np.pi / 3.0 * float(kg) ** 3
    assert abs(knorm - 25.4921722) < 1e-5
    assert abs(knorm * 1.23485 ** 3 - 48.0009) < 1e-3
    # uniform lattice: rho ~ 1 (analytic solution of tests/test125cells.c:148-207)
    ic = ics.lattice(12, jitter=0.0, sedov=False)
    prm = ol.make_params()
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    o = ol.run_step("oracle", prm, ics.permute(ic, order), cells, tasks, nt)
    assert np.abs(o["g_rho"] - 1.0).max() < 2e-2
    assert np.abs(o["f_a"]).max() < 1e-3 * float(o["g_pressure"].max()) / float(ic["h"][0])


def test_force_pairs_symmetric():
    """With every particle active the force loop is symmetric: (i,j) is applied iff (j,i) is
    (tests/testSymmetry.c), and the density pairs are a subset of the force pairs when h is converged."""
    ic = ics.uniform_random(3000, seed=8)
    prm = ol.make_params(record=True)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    o = ol.run_step("oracle", prm, ics.permute(ic, order), cells, tasks, nt, record=True)
    p = o["pairs"][2]
    fwd = set(map(tuple, p.tolist()))
    assert len(fwd) == len(p), "no ordered pair is applied twice"
    assert all((j, i) in fwd for (i, j) in fwd)


def test_threaded_reference_harness_agrees_with_the_serial_one():
    """The timed CPU arm of bench.py shares ONE box among pthreads (oracle/ref_harness.c, ref_set_threads): shared
    task counter, top-level cell locks.  Sums then arrive in another order, so the comparison with the serial run
    (the form every fixture is made and checked with) is within the h tolerance of the ghost, not bit for bit."""
    import ctypes
    if not ol.ref_available("minimal"):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    ic = ics.uniform_random(24 ** 3, seed=17)
    prm = ol.make_params()
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    icc = ics.permute(ic, order)
    lib = ol.ref("minimal")
    lib.ref_set_threads.argtypes = [ctypes.c_int]
    lib.ref_last_step_seconds.restype = ctypes.c_double
    serial = ol.run_step("ref", prm, icc, cells, tasks, nt)
    try:
        lib.ref_set_threads(4)
        threaded = ol.run_step("ref", prm, icc, cells, tasks, nt)
        assert lib.ref_last_step_seconds() > 0.0
    finally:
        lib.ref_set_threads(1)
    assert threaded["counts"] == [0, 0, 0], "nothing is counted or recorded in the threaded mode"
    for k, tol in (("g_h", 2e-4), ("g_rho", 2e-4), ("f_a", 1e-3), ("f_u_dt", 1e-3)):
        a, b = threaded[k].astype(np.float64), serial[k].astype(np.float64)
        assert np.abs(a - b).max() <= tol * np.abs(b).max(), k
    again = ol.run_step("ref", prm, icc, cells, tasks, nt)
    for k in FIELDS:
        assert np.array_equal(again[k], serial[k]), "the serial mode is what it was: " + k


def _brute_force_counts(x, H, active, dim, periodic):
    """All pairs, no tree: for every active i the neighbours within its own kernel support (what the gradient loop
    of SPHENIX and a density pass with the converged h visit) and within either support (the force loop)."""
    n = len(H)
    H2 = H.astype(np.float32) ** 2
    n_own, n_either = np.zeros(n, np.int64), np.zeros(n, np.int64)
    for i in np.nonzero(active)[0]:
        dx = x[i] - x
        if periodic:
            dx -= np.round(dx / dim) * dim
        r2 = (dx.astype(np.float32) ** 2).sum(1)
        own = r2 < H2[i]
        own[i] = False
        either = own | (r2 < H2)
        either[i] = False
        n_own[i], n_either[i] = own.sum(), either.sum()
    return n_own, n_either


def _exact_count_inputs(g):
    kw = g["params"]
    tb = g["ic"].get("time_bin")
    act = np.ones(len(g["h"]), bool) if tb is None else tb <= kw.get("max_active_bin", 56)
    gamma = np.float32(1.825742)
    # active particles: the converged h; the others keep theirs
    H = np.where(act, g["g_h"] * gamma, g["h"] * gamma).astype(np.float32)
    return H, act, np.array(kw.get("dim", (1.0, 1.0, 1.0)), dtype=np.float64), kw.get("periodic", True)


@pytest.mark.parametrize("name", gu.names())
def test_force_loop_counts_against_brute_force(name):
    """The role of the reference's exact-count checker (hydro_exact_density_compute / N_force_exact,
    src/hydro.c:50-260 under SWIFT_HYDRO_DENSITY_CHECKS): the number of neighbours every particle met in the
    reference's force loop, from positions and smoothing lengths alone."""
    g = gu.load(name)
    H, act, dim, periodic = _exact_count_inputs(g)
    n_own, n_either = _brute_force_counts(g["x"], H, act, dim, periodic)
    assert np.array_equal(n_either, g["ngb2"])
    assert int(n_either.sum()) == int(g["counts"][2])
