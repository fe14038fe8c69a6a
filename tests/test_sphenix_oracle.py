"""CPU tests of the checker for the three-loop scheme (SPHENIX): the plain-C oracle against
the golden vectors written by the reference's own compiled SPHENIX operators (density, ghost with
hydro_prepare_gradient, the gradient loop, runner_do_extra_ghost, force), and live against the compiled
reference when oracle/_ref is present."""
import numpy as np
import pytest

import golden_util as gu
import oracle_lib as ol
from swift_subtask_dev_b200 import ics

FIELDS = list(ol.OUT_FIELDS) + list(ol.SPHENIX_OUT) + ["f_min_ngb_time_bin"]


def _active(g):
    if g.get("time_bin") is None:
        return np.ones(len(g["h"]), dtype=bool)
    return g["time_bin"] <= g["params"].get("max_active_bin", 56)


@pytest.mark.parametrize("name", gu.sphenix_names())
def test_oracle_reproduces_reference_golden_sphenix(name):
    g = gu.load_sphenix(name)
    cells, tasks, nt = gu.tree_structs(g)
    prm = ol.make_params(record=True, scheme=2, **g["params"])
    o = ol.run_step("oracle", prm, g["ic"], cells, tasks, nt, record=True, sphenix=g["props"])
    assert o["counts"] == [int(v) for v in g["counts"]]
    for s in range(4):  # 0 density, 1 ghost redo, 2 force, 3 gradient
        assert gu.pair_digest(o["pairs"][s]) == str(g["digests"][s]), "pair list of loop %d differs" % s
        assert np.array_equal(gu.ngb_counts(o["pairs"][s], g["id"]), g["ngb%d" % s])
    act = _active(g)
    for k in FIELDS:
        assert np.array_equal(o[k][act], g[k][act]), "field %s is not bit-identical to the reference's" % k
    # the fixtures exercise what they are for
    assert np.count_nonzero(g["e_alpha_visc"][act] != g["in_alpha_visc"][act]) > 0.9 * act.sum()
    assert np.abs(g["e_laplace_u"][act]).max() > 0 and g["counts"][3] > 0


def test_three_fixtures_present():
    assert len(gu.sphenix_names()) >= 3


LIVE = [
    ("uniform_split", lambda: ics.uniform_random(30000, seed=4), dict(periodic=True)),
    # (seed chosen so that no two particles of a cell that share an f32 sort key -- their order is unspecified in
    # the reference's quicksort -- feed the same particle's sums: every field is then bit-identical)
    ("clustered", lambda: ics.clustered(12000, nclumps=6, seed=7), dict(periodic=False)),
]


@pytest.mark.parametrize("name,make,kw", LIVE, ids=[c[0] for c in LIVE])
def test_oracle_vs_compiled_reference_sphenix(name, make, kw):
    if not ol.ref_available("sphenix"):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    ic = make()
    rng = np.random.default_rng(11)
    n = len(ic["h"])
    ic["u"] = (ic["u"] * (1.0 + 0.5 * rng.random(n))).astype(np.float32)
    ic["in_alpha_visc"] = (0.1 + rng.random(n)).astype(np.float32)
    ic["in_alpha_diff"] = (0.5 * rng.random(n)).astype(np.float32)
    ic["in_div_v_prev"] = rng.standard_normal(n).astype(np.float32)
    prm = ol.make_params(record=True, scheme=2, alpha=0.1, **kw)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
    icc = ics.permute(ic, order)
    o = ol.run_step("oracle", prm, icc, cells, tasks, nt, record=True, sphenix=ol.SPHENIX_PROPS)
    r = ol.run_step("ref", prm, icc, cells, tasks, nt, record=True, sphenix=ol.SPHENIX_PROPS)
    assert o["counts"] == r["counts"]
    for s in range(4):
        assert gu.pair_digest(o["pairs"][s]) == gu.pair_digest(r["pairs"][s])
    for k in FIELDS:
        assert np.array_equal(o[k], r[k]), k


# ---- engine steps with individual time bins (drift, the seven hydro phases, kick2, time-step, kick1) ----
def _integ_case():
    import importlib.util
    import os
    path = os.path.join(gu.GOLDEN_DIR, "make_golden_sphenix.py")
    spec = importlib.util.spec_from_file_location("make_golden_sphenix", path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    icc, tree, kw, ip, nsteps = m.integ_case()
    z = np.load(os.path.join(gu.SPHENIX_DIR, "integ", "clustered4k_bins.npz"))
    return icc, tree, kw, ip, nsteps, z, m.INTEG_KEYS


def test_oracle_engine_loop_reproduces_reference_golden_sphenix():
    """The state the reference's own cell_drift_part / runner_do_kick1,2 / runner_do_timestep and its SPHENIX hydro
    operators leave after six engine steps (tests/golden/sphenix/integ), alphas included: identical bits."""
    import integ_lib as il
    icc, tree, kw, ip, nsteps, z, extra = _integ_case()
    s, log = il.run_loop("oracle", icc, kw, ip, tree, nsteps, sphenix=ol.SPHENIX_PROPS)
    for k in il.STATE_KEYS + extra:
        assert np.array_equal(np.asarray(s[k]).reshape(-1), z[k].reshape(-1)), k
    assert [[l[0], l[1], l[2], l[5]] for l in log] == z["log_int"].tolist()
    assert len(np.unique(s["time_bin"])) >= 4, "not a time-bin hierarchy"
    assert s["alpha_visc"].max() > 1.0, "the viscosity switch should have fired somewhere"


@pytest.mark.parametrize("name", gu.sphenix_names())
def test_gradient_and_force_loop_counts_against_brute_force(name):
    """N_gradient_exact and N_force_exact of the reference's checker (src/hydro.c:50-260, SWIFT_HYDRO_DENSITY_CHECKS)
    from positions and converged smoothing lengths alone, all pairs: the gradient loop visits the neighbours within a
    particle's own kernel support, the force loop those within either support."""
    from test_oracle import _brute_force_counts, _exact_count_inputs
    g = gu.load_sphenix(name)
    H, act, dim, periodic = _exact_count_inputs(g)
    n_own, n_either = _brute_force_counts(g["x"], H, act, dim, periodic)
    assert np.array_equal(n_own, g["ngb3"]), "gradient loop"
    assert np.array_equal(n_either, g["ngb2"]), "force loop"
    assert int(n_own.sum()) == int(g["counts"][3]) and int(n_either.sum()) == int(g["counts"][2])
