"""GPU parity tests of the three-loop scheme (SPHENIX): the CUDA path through the C ABI --
density, ghost with hydro_prepare_gradient, the GRADIENT loop (k_density<0, false, 1>), runner_do_extra_ghost
(k_extra_ghost), the scheme's force loop (k_force<2>) -- against the golden vectors the reference's own compiled
SPHENIX operators wrote (tests/golden/sphenix) and against the oracle on a library-built tree.
Bit-exact: counts, the ordered pair sets of all FOUR loops, every field after every stage."""
import ctypes as C

import numpy as np
import pytest

import golden_util as gu
import oracle_lib as ol
import parity
from swift_subtask_dev_b200 import HydroBackend, ics
from swift_subtask_dev_b200 import backend as B

pytestmark = pytest.mark.gpu

# GPU key -> reference key, by the stage after which the reference value was taken
AFTER_GHOST = [("h", "g_h"), ("rho", "g_rho"), ("f", "g_f"), ("pressure", "g_pressure"), ("soundspeed", "g_soundspeed"),
               ("balsara", "g_balsara"), ("d_rho", "d_rho"), ("d_rho_dh", "d_rho_dh"), ("d_wcount", "d_wcount"),
               ("d_wcount_dh", "d_wcount_dh"), ("d_div_v", "d_div_v"), ("d_rot_v", "d_rot_v"), ("div_v", "g_div_v")]
AFTER_GRADIENT = [("v_sig", "r_v_sig"), ("laplace_u", "r_laplace_u"), ("alpha_visc_max_ngb", "r_alpha_visc_max_ngb")]
AFTER_EXTRA_GHOST = [("laplace_u", "e_laplace_u"), ("alpha_visc", "e_alpha_visc"), ("alpha_diff", "e_alpha_diff"),
                     ("div_v_previous_step", "e_div_v_prev"), ("div_v_dt", "e_div_v_dt")]
AFTER_FORCE = [("a", "f_a"), ("u_dt", "f_u_dt"), ("h_dt", "f_h_dt"), ("v_sig", "f_v_sig"),
               ("min_ngb_time_bin", "f_min_ngb_time_bin")]


def _backend(kw, props, fields_mode=0):
    prm = parity.gpu_params(dict(kw, scheme=B.SCHEME_SPHENIX))
    prm.fields_mode = fields_mode
    be = HydroBackend(prm)
    be.sphenix_set_props(**props)
    return be


def _set_state(be, ic):
    be.sphenix_set_state(ic.get("in_alpha_visc"), ic.get("in_alpha_diff"), ic.get("in_div_v_prev"))


def _both(be):
    out = be.results(first_pass=True)
    out.update(be.sphenix_results())
    return out


def _assert_fields(got, ref, table, act, stage):
    for kg, ko in table:
        a, b = np.asarray(got[kg]), np.asarray(ref[ko])
        if act is not None:
            a, b = a[act], b[act]
        bad = a != b
        if a.dtype.kind == "f":
            bad &= ~(np.isnan(a) & np.isnan(b))
        n = int(np.count_nonzero(bad.reshape(len(a), -1).any(axis=1))) if a.size else 0
        assert n == 0, "after %s: field %s differs from the reference's %s on %d particles (max abs %g)" % (
            stage, kg, ko, n, float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64)))))


def _active(g):
    if g["ic"].get("time_bin") is None:
        return None
    return g["ic"]["time_bin"] <= g["params"].get("max_active_bin", 56)


def _build_golden(g, record, fields_mode=0):
    cells, tasks, nt = gu.tree_structs(g)
    be = _backend(g["params"], g["props"], fields_mode)
    gc = (B.Cell * len(cells))()
    C.memmove(gc, cells, C.sizeof(B.Cell) * len(cells))
    gt = (B.Task * max(nt, 1))()
    C.memmove(gt, tasks, C.sizeof(B.Task) * nt)
    be.build_from_cells(g["ic"], gc, gt, nt)
    _set_state(be, g["ic"])
    if record:
        be.record_pairs(len(g["h"]) * 400)
    return be


@pytest.mark.parametrize("name", gu.sphenix_names())
def test_sphenix_golden_phase_by_phase(name):
    """The operator-level entries in the reference's task order, every field compared after every stage."""
    g = gu.load_sphenix(name)
    act = _active(g)
    n = len(g["h"])
    be = _build_golden(g, record=True)
    be.runner_do_hydro_sort()
    be.runner_do_density()
    be.runner_do_ghost()
    _assert_fields(_both(be), g, AFTER_GHOST, act, "ghost")
    be.runner_do_gradient()
    _assert_fields(_both(be), g, AFTER_GRADIENT, act, "gradient loop")
    be.runner_do_extra_ghost()
    _assert_fields(_both(be), g, AFTER_EXTRA_GHOST, act, "extra ghost")
    be.runner_do_force()
    out = _both(be)
    _assert_fields(out, g, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST[:6], act, "force")
    c = be.counters()
    assert [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]] == [int(v) for v in g["counts"]]
    for s in range(4):  # 0 density, 1 ghost redo, 2 force, 3 gradient
        p = be.fetch_pairs(s, n * 400)
        assert gu.pair_digest(p) == str(g["digests"][s]), "pair list of loop %d differs" % s
        assert np.array_equal(gu.ngb_counts(p, g["id"]), g["ngb%d" % s])
    be.close()


@pytest.mark.parametrize("name", gu.sphenix_names())
def test_sphenix_golden_whole_step(name):
    """swift_cuda_step runs the seven phases of the scheme in one call: the same final state."""
    g = gu.load_sphenix(name)
    act = _active(g)
    be = _build_golden(g, record=False)
    be.step()
    out = _both(be)
    _assert_fields(out, g, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST, act, "step")
    c = be.counters()
    assert [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]] == [int(v) for v in g["counts"]]
    be.close()


@pytest.mark.parametrize("name", gu.sphenix_names())
def test_sphenix_golden_fast_fields_mode(name):
    """fields_mode = FAST (k_force_fast<2>: FMA / MUFU arithmetic in the scheme's force interaction): the pair lists
    of all four loops, the counts and every field up to and including the extra ghost bit-exact; a, u_dt, h_dt
    within 1e-4 (BASELINE.json: north_star) relative to max(|reference|, 0.1 x the rms of the field)."""
    g = gu.load_sphenix(name)
    act = _active(g)
    n = len(g["h"])
    be = _build_golden(g, record=True, fields_mode=1)
    be.step()
    out = _both(be)
    _assert_fields(out, g, AFTER_EXTRA_GHOST + AFTER_GHOST + [("v_sig", "f_v_sig"), ("min_ngb_time_bin", "f_min_ngb_time_bin")],
                   act, "step (FAST)")
    c = be.counters()
    assert [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]] == [int(v) for v in g["counts"]]
    for s in range(4):
        assert gu.pair_digest(be.fetch_pairs(s, n * 400)) == str(g["digests"][s]), "pair list of loop %d differs" % s
    be.close()
    worst = {}
    for kg, ko in (("a", "f_a"), ("u_dt", "f_u_dt"), ("h_dt", "f_h_dt")):
        a, b = np.asarray(out[kg], dtype=np.float64), np.asarray(g[ko], dtype=np.float64)
        if act is not None:
            a, b = a[act], b[act]
        rms = float(np.sqrt(np.mean(b ** 2)))
        worst[kg] = float((np.abs(a - b) / np.maximum(np.abs(b), max(0.1 * rms, 1e-30))).max())
        assert worst[kg] <= parity.TOL_FORCE, "FAST field %s: relative error %g" % (kg, worst[kg])
    print(name, "SPHENIX FAST worst relative errors:", worst)


def _oracle_on_library_tree(ic, tree, kw, record=False):
    """The oracle on the library's own cells and tasks: inputs permuted into its cell order, outputs back."""
    n = len(ic["h"])
    order = tree["order"]
    icc = ics.permute(ic, order)
    cells = (ol.Cell * tree["ncells"])()
    C.memmove(cells, tree["cells"], C.sizeof(ol.Cell) * tree["ncells"])
    tasks = (ol.Task * max(tree["ntasks"], 1))()
    C.memmove(tasks, tree["tasks"], C.sizeof(ol.Task) * tree["ntasks"])
    prm = ol.make_params(record=record, scheme=2, **kw)
    o = ol.run_step("oracle", prm, icc, cells, tasks, tree["ntasks"], record=record, sphenix=ol.SPHENIX_PROPS)
    inv = np.empty_like(order)
    inv[order] = np.arange(len(order), dtype=order.dtype)
    for k in list(o.keys()):
        if isinstance(o[k], np.ndarray) and len(o[k]) == n:
            o[k] = o[k][inv]
    return o


def _state(ic, seed):
    rng = np.random.default_rng(seed)
    n = len(ic["h"])
    ic["u"] = (ic["u"] * (1.0 + 0.5 * rng.random(n))).astype(np.float32)
    ic["in_alpha_visc"] = (0.1 + rng.random(n)).astype(np.float32)
    ic["in_alpha_diff"] = (0.5 * rng.random(n)).astype(np.float32)
    ic["in_div_v_prev"] = rng.standard_normal(n).astype(np.float32)
    return ic


def test_sphenix_library_tree_clustered_pairs():
    """Library-built tree over a clumpy box (split cells, tasks pushed down to the progeny): the oracle runs on
    the library's own tree and task list; pair SETS of the four loops and every field identical."""
    ic = _state(ics.clustered(20000, nclumps=8, seed=6), 11)
    n = len(ic["h"])
    kw = dict(periodic=True, alpha=0.1)
    be = _backend(kw, ol.SPHENIX_PROPS)
    be.build(ic)
    _set_state(be, ic)
    be.record_pairs(n * 600)
    be.step()
    out = _both(be)
    c = be.counters()
    tree = be.tree()
    pairs = [be.fetch_pairs(s, n * 600) for s in range(4)]
    be.close()
    assert any(tree["cells"][k].split for k in range(tree["ncells"])), "the case should have split cells"
    o = _oracle_on_library_tree(ic, tree, kw, record=True)
    assert [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]] == o["counts"]
    for s in range(4):
        assert parity.pair_set_diff(pairs[s], o["pairs"][s]) == 0, "pair set of loop %d differs" % s
    _assert_fields(out, o, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST, None, "step")


def test_sphenix_library_tree_sub_tasks():
    """96^3 in 12^3 top-level cells: 512 particles per cell, split once, sub_self / sub_pair tasks that the library
    flattens to leaf interactions and the oracle recurses through: counts and every field identical."""
    ic = _state(ics.uniform_random(96 ** 3, seed=5), 12)
    kw = dict(periodic=True, alpha=0.1)
    be = _backend(kw, ol.SPHENIX_PROPS)
    be.build(ic)
    _set_state(be, ic)
    be.step()
    out = _both(be)
    c = be.counters()
    tree = be.tree()
    be.close()
    assert {2, 3} <= {tree["tasks"][k].type for k in range(tree["ntasks"])}, "the case should have sub tasks"
    o = _oracle_on_library_tree(ic, tree, kw)
    assert [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]] == o["counts"]
    _assert_fields(out, o, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST, None, "step")


def test_sphenix_state_is_carried_on_the_device():
    """Two steps in a row: the viscosity / diffusion alphas and div_v_previous_step the first step leaves stay in
    HBM and are what the second step starts from (only x, v, h, u go up again), as p->viscosity / p->diffusion
    persist in the reference's struct part.  The oracle gets the first step's outputs as the second one's inputs.
    (Converged smoothing lengths from the start, as after the reference's engine_init_particles: the top-level
    grid is then sized for them.)"""
    import integ_lib as il
    kw = dict(periodic=True, alpha=0.1, scheme=2)
    ic = ics.uniform_random(16 ** 3, seed=8)
    ic["v"] = (ic["v"] - 0.3 * (ic["x"] - 0.5)).astype(np.float32)
    icc, (cells, tasks, nt) = il.prepare(ic, kw)
    icc = _state(icc, 13)
    be = _backend(kw, ol.SPHENIX_PROPS)
    gc = (B.Cell * len(cells))()
    C.memmove(gc, cells, C.sizeof(B.Cell) * len(cells))
    gt = (B.Task * max(nt, 1))()
    C.memmove(gt, tasks, C.sizeof(B.Task) * nt)
    be.build_from_cells(icc, gc, gt, nt)
    _set_state(be, icc)
    be.step()
    r1 = _both(be)
    ic2 = dict(icc)
    ic2["h"] = r1["h"].copy()
    be.update_particles(ic2)
    be.step()
    r2 = _both(be)
    c2 = be.counters()
    be.close()
    prm = ol.make_params(**kw)
    o1 = ol.run_step("oracle", prm, icc, cells, tasks, nt, sphenix=ol.SPHENIX_PROPS)
    _assert_fields(r1, o1, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST, None, "step 1")
    ic2["in_alpha_visc"], ic2["in_alpha_diff"], ic2["in_div_v_prev"] = o1["e_alpha_visc"], o1["e_alpha_diff"], \
        o1["e_div_v_prev"]
    o2 = ol.run_step("oracle", prm, ic2, cells, tasks, nt, sphenix=ol.SPHENIX_PROPS)
    assert [c2["n_density"], c2["n_density_redo"], c2["n_force"], r2["n_gradient"]] == o2["counts"]
    _assert_fields(r2, o2, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST, None, "step 2")
    assert not np.array_equal(r1["alpha_visc"], r2["alpha_visc"]), "the alphas should have moved between the steps"


def test_sphenix_config0_lattice64():
    """BASELINE.json configs[0]: SedovBlast_3D-style jittered lattice, 64^3 particles, the DEFAULT SPH scheme (which is
    SPHENIX, configure.ac:1958-1959), one step: every field of every particle against the oracle, bit for bit (the
    oracle orders equal sort keys by index, as k_sort does; the reference's quicksort leaves that order unspecified and
    differs from both on a handful of lattice particles by an ulp)."""
    ic = ics.lattice(64, jitter=0.1, seed=1234)
    kw = dict(periodic=True, alpha=0.1)
    n = len(ic["h"])
    be = _backend(kw, ol.SPHENIX_PROPS)
    be.build(ic)
    be.step()
    out = _both(be)
    c = be.counters()
    tree = be.tree()
    be.close()
    o = _oracle_on_library_tree(ic, tree, kw)
    assert [c["n_density"], c["n_density_redo"], c["n_force"], out["n_gradient"]] == o["counts"]
    _assert_fields(out, o, AFTER_FORCE + AFTER_EXTRA_GHOST + AFTER_GHOST, None, "step")
    # the Sedov energy spike is what the scheme sees in this configuration (the gas is at rest): del^2 u
    assert np.abs(out["laplace_u"]).max() > 0 and out["alpha_diff"].max() > 0


def test_sphenix_resident_engine_loop():
    """N3 with the three-loop scheme: the particles go to the device once; drift, the seven hydro phases on the
    active particles, kick2, time-step, kick1 per engine step with individual time bins.  Every field -- the
    viscosity / diffusion alphas, div_v_previous_step, div_v_dt, laplace_u included -- against the oracle after
    EVERY step and against the reference's golden state at the end."""
    import integ_lib as il
    import test_sphenix_oracle as cpu
    icc, tree, kw, ip, nsteps, z, extra = cpu._integ_case()
    keys = il.STATE_KEYS + extra
    ref_states = []
    so, lo = il.run_loop("oracle", icc, kw, ip, tree, nsteps, sphenix=ol.SPHENIX_PROPS,
                         on_step=lambda step, s: ref_states.append({k: s[k].copy() for k in keys}))
    be = _backend(dict(kw, max_active_bin=il.NUM_TIME_BINS), ol.SPHENIX_PROPS)

    def check(step, g):
        for k in keys:
            assert np.array_equal(np.asarray(g[k]).reshape(-1), ref_states[step][k].reshape(-1)), \
                "step %d: field %s is not bit-identical" % (step, k)

    sg, lg = il.run_loop_gpu(be, icc, ip, tree, nsteps, on_step=check, sphenix=True)
    be.close()
    assert [l[:3] + l[5:] for l in lg] == [l[:3] + l[5:] for l in lo], "active counts / sync points differ"
    for k in keys:
        assert np.array_equal(np.asarray(sg[k]).reshape(-1), z[k].reshape(-1)), "%s vs the reference's golden state" % k


# ---- struct part in, struct part out: the scheme's 160-byte record (src/hydro/SPHENIX/hydro_part.h) ----
REC_AFTER_DENSITY = [("rho", "d_rho"), ("rho_dh", "d_rho_dh"), ("wcount", "d_wcount"), ("wcount_dh", "d_wcount_dh"),
                     ("div_v", "d_div_v"), ("rot_v", "d_rot_v")]
REC_AFTER_GHOST = [("h", "g_h"), ("rho", "g_rho"), ("f", "g_f"), ("pressure", "g_pressure"), ("soundspeed", "g_soundspeed"),
                   ("balsara", "g_balsara"), ("div_v", "g_div_v")]
REC_AFTER_GRADIENT = [("v_sig", "r_v_sig"), ("laplace_u", "r_laplace_u"), ("alpha_visc_max_ngb", "r_alpha_visc_max_ngb")]
REC_AFTER_EXTRA = [("laplace_u", "e_laplace_u"), ("alpha_visc", "e_alpha_visc"), ("alpha_diff", "e_alpha_diff"),
                   ("div_v_previous_step", "e_div_v_prev"), ("div_v_dt", "e_div_v_dt")]
REC_AFTER_FORCE = [("a_hydro", "f_a"), ("u_dt", "f_u_dt"), ("h_dt", "f_h_dt"), ("v_sig", "f_v_sig"),
                   ("min_ngb_time_bin", "f_min_ngb_time_bin")]


def _records(g):
    """The caller's array of struct part as it is before the step: inputs, the carried state of the scheme, and
    for inactive particles what their last active step left in the force union."""
    parts = B.parts_from_ic(g["ic"], scheme=B.SCHEME_SPHENIX)
    parts["min_ngb_time_bin"] = 57   # num_time_bins + 1, as the reference's time-step task leaves it
    for rec, key in (("rho", "in_rho"), ("pressure", "in_pressure"), ("f", "in_f"), ("soundspeed", "in_soundspeed"),
                     ("balsara", "in_balsara")):
        if g["ic"].get(key) is not None:
            parts[rec] = g["ic"][key]
    return parts


def _parts_backend(g, parts):
    cells, tasks, nt = gu.tree_structs(g, B.Cell, B.Task)
    be = _backend(g["params"], g["props"])
    be.build_from_parts(parts, cells, tasks, nt)
    return be


@pytest.mark.parametrize("name", gu.sphenix_names())
def test_sphenix_struct_part_step(name):
    """swift_cuda_step_parts on the scheme's records, all seven phases in one call: every field the reference's
    operators write, in the record, bit for bit; inactive records untouched."""
    g = gu.load_sphenix(name)
    act = _active(g)
    parts = _records(g)
    before = parts.copy()
    be = _parts_backend(g, parts)
    be.step_parts(parts, B.PHASE_ALL_SPHENIX)
    c = be.counters()
    ngrad = be.sphenix_results()["n_gradient"]
    be.close()
    assert [c["n_density"], c["n_density_redo"], c["n_force"], ngrad] == [int(v) for v in g["counts"]]
    table = REC_AFTER_GHOST + REC_AFTER_EXTRA + REC_AFTER_FORCE + [("alpha_visc_max_ngb", "r_alpha_visc_max_ngb")]
    _assert_fields(parts, g, table, act, "step_parts")
    if act is not None:
        for k in parts.dtype.names:  # (field by field: the gaps of the record dtype are not copied by numpy)
            assert np.array_equal(parts[k][~act], before[k][~act]), "field %s of an inactive record was written" % k


@pytest.mark.parametrize("name", ["uniform14_periodic", "random11_timebins"])
def test_sphenix_struct_part_phase_calls(name):
    """One call per phase on the records, as a per-task shim makes them (each call uploads the records as they are
    and writes back what its phase produces): the record after every call against the reference's stage values."""
    g = gu.load_sphenix(name)
    act = _active(g)
    parts = _records(g)
    be = _parts_backend(g, parts)
    be.step_parts(parts, B.PHASE_SORT | B.PHASE_DENSITY)
    _assert_fields(parts, g, REC_AFTER_DENSITY, act, "density call")
    be.step_parts(parts, B.PHASE_GHOST)
    _assert_fields(parts, g, REC_AFTER_GHOST, act, "ghost call")
    be.step_parts(parts, B.PHASE_GRADIENT)
    _assert_fields(parts, g, REC_AFTER_GRADIENT + REC_AFTER_GHOST, act, "gradient call")
    be.step_parts(parts, B.PHASE_EXTRA_GHOST)
    _assert_fields(parts, g, REC_AFTER_EXTRA + REC_AFTER_GHOST, act, "extra ghost call")
    be.step_parts(parts, B.PHASE_FORCE | B.PHASE_END_FORCE)
    be.close()
    _assert_fields(parts, g, REC_AFTER_FORCE + REC_AFTER_EXTRA + REC_AFTER_GHOST, act, "force call")


@pytest.mark.parametrize("nranks", [2, 4])
def test_sphenix_ranks_reproduce_single_gpu_run(nranks):
    """Several ranks (all in this process, halos moved device to device) with the scheme's THREE exchanges per step --
    xv; rho + u + the viscosity alpha of before the extra ghost; the evolved alphas after it -- against one context
    holding the whole box: every owned particle's every field, the scheme's included, bit-identical."""
    from swift_subtask_dev_b200 import multi
    grid = multi.grid_for(nranks)
    dim = tuple(float(g) for g in grid)
    per_cell = 400 if nranks == 2 else 150
    cdim = tuple(int(12 * d / max(dim)) for d in dim)
    ic = ics.cell_block_random(np.arange(int(np.prod(cdim))), cdim, dim, per_cell, seed=7)
    ident = ic["id"].astype(np.int64)
    ic["u"] = (1.0 + 0.3 * (ident % 13) / 13.0).astype(np.float32)
    kw = dict(scheme=B.SCHEME_SPHENIX, alpha=0.1)

    def carried(ids):
        return ((0.1 + 0.05 * (ids % 7)).astype(np.float32), (0.02 * (ids % 5)).astype(np.float32),
                (0.01 * ((ids % 11) - 5)).astype(np.float32))

    part, states = multi.setup_local(nranks, dim, cdim, per_cell, seed=7, params_kw=kw, ic=ic)
    for st in states:
        st.be.sphenix_set_state(*carried(st.ic["id"].astype(np.int64)))
    multi.step_ranks_local(states)
    be = HydroBackend(B.make_params(dim=dim, periodic=True, **kw))
    be.build(ic)
    be.sphenix_set_state(*carried(ident))
    be.step()
    ref = _both(be)
    cref = be.counters()
    be.close()
    pos = {int(v): k for k, v in enumerate(ic["id"])}
    tot = np.zeros(4, dtype=np.int64)
    for st in states:
        r = st.owned_results()
        idx = np.array([pos[int(v)] for v in r["id"]])
        for k in r:
            if k != "id":
                assert np.array_equal(r[k], ref[k][idx]), (st.rank, k)
        c = st.be.counters()
        tot += np.array([c["n_density"], c["n_density_redo"], c["n_force"], st.be.sphenix_results()["n_gradient"]])
        assert st.halo_bytes_per_step() > 0
        st.be.close()
    assert tuple(tot) == (cref["n_density"], cref["n_density_redo"], cref["n_force"], ref["n_gradient"])


def test_sphenix_step_io_and_refusals():
    """swift_cuda_step_io (upload, the seven phases, download overlapped) gives the bits of the separate calls; the
    two phases of the scheme are refused in a two-loop context."""
    ic = _state(ics.uniform_random(20000, seed=21), 3)
    kw = dict(periodic=True, alpha=0.1)
    be = _backend(kw, ol.SPHENIX_PROPS)
    be.build(ic)
    _set_state(be, ic)
    be.update_particles(ic)
    be.step()
    a = _both(be)
    _set_state(be, ic)   # the first step evolved the alphas
    out = be.alloc_results()
    be.step_io(ic, out)
    b = be.sphenix_results()
    be.close()
    for k in out:
        assert np.array_equal(a[k], out[k]), k
    for k in b:
        assert np.array_equal(a[k], b[k]), k
    ic = ics.uniform_random(1000, seed=1)
    # and the two new phases exist in that scheme only
    be = HydroBackend(parity.gpu_params(dict(periodic=True)))
    be.build(ic)
    with pytest.raises(B.SwiftCudaError, match="SPHENIX scheme only"):
        be.runner_do_gradient()
    be.close()
