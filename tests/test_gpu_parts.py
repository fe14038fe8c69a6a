"""struct part in, struct part out: the AoS entries of the C ABI
(swift_cuda_build_from_parts, swift_cuda_step_parts) against the SoA entries on the same particles:
identical bits in every field, and the per-task phase calls the C shim makes (density with the
running sums kept, ghost, force, end_force) against the whole step."""
import numpy as np
import pytest

from swift_subtask_dev_b200 import HydroBackend, ics, make_params
from swift_subtask_dev_b200 import backend as be


def soa_step(ic, kw):
    b = HydroBackend(make_params(**kw))
    b.build(ic)
    b.step()
    out = b.results(first_pass=True)
    cnt = b.counters()
    b.close()
    return out, cnt


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["uniform", "clustered", "timebins"])
def test_step_parts_matches_the_soa_step(case):
    if case == "uniform":
        ic = ics.uniform_random(30 ** 3, seed=3)
    elif case == "clustered":
        ic = ics.clustered(20000, nclumps=8, seed=4)
    else:
        ic = ics.uniform_random(24 ** 3, seed=5)
        rng = np.random.default_rng(1)
        ic["time_bin"] = rng.integers(1, 4, size=len(ic["h"])).astype(np.int8)
    kw = dict(periodic=True, max_active_bin=2 if case == "timebins" else 56)
    ref, cref = soa_step(ic, kw)
    parts = be.parts_from_ic(ic)
    if case == "timebins":
        # what inactive particles carry from their last active step
        for k in ("rho", "pressure", "soundspeed"):
            parts[k] = 1.0
        ic = dict(ic, in_rho=parts["rho"].copy(), in_pressure=parts["pressure"].copy(),
                  in_soundspeed=parts["soundspeed"].copy())
        ref, cref = soa_step(ic, kw)
    b = HydroBackend(make_params(**kw))
    b.build_from_parts(parts)
    b.step_parts(parts)
    cnt = b.counters()
    b.close()
    for k in ("n_density", "n_density_redo", "n_force"):
        assert cnt[k] == cref[k], k
    act = parts["time_bin"] <= kw["max_active_bin"]
    for rec, key in (("h", "h"), ("rho", "rho"), ("pressure", "pressure"), ("soundspeed", "soundspeed"), ("f", "f"),
                     ("balsara", "balsara"), ("a_hydro", "a"), ("u_dt", "u_dt"), ("h_dt", "h_dt"), ("v_sig", "v_sig"),
                     ("min_ngb_time_bin", "min_ngb_time_bin")):
        assert np.array_equal(parts[rec][act], ref[key][act]), rec
    # inactive records are untouched
    assert (parts["a_hydro"][~act] == 0).all()


@pytest.mark.gpu
def test_phase_calls_add_up_to_the_step():
    """density (two halves of the task list are not separable here, so: density alone), then ghost on the
    stored sums, then force, then end_force, each a separate call on the records: the same bits as one call."""
    ic = ics.uniform_random(26 ** 3, seed=8)
    kw = dict(periodic=True)
    whole = be.parts_from_ic(ic)
    b = HydroBackend(make_params(**kw))
    b.build_from_parts(whole)
    b.step_parts(whole)
    parts = be.parts_from_ic(ic)
    b.step_parts(parts, be.PHASE_SORT | be.PHASE_DENSITY)
    first = parts.copy()
    assert (first["wcount"] > 0).all()
    b.step_parts(parts, be.PHASE_GHOST)
    b.step_parts(parts, be.PHASE_FORCE)
    b.step_parts(parts, be.PHASE_END_FORCE)
    b.close()
    for k in ("h", "rho", "pressure", "soundspeed", "f", "balsara", "a_hydro", "u_dt", "h_dt", "v_sig"):
        assert np.array_equal(parts[k], whole[k]), k
    # KEEP: a second density call continues the sums of the first (here: doubles every term's count)
    b = HydroBackend(make_params(**kw))
    twice = be.parts_from_ic(ic)
    b.build_from_parts(twice)
    b.step_parts(twice, be.PHASE_SORT | be.PHASE_DENSITY)
    once = twice["wcount"].copy()
    b.step_parts(twice, be.PHASE_DENSITY | be.PHASE_KEEP)
    b.close()
    assert np.allclose(twice["wcount"], 2 * once, rtol=1e-5)
    assert (twice["wcount"] > once).all()


@pytest.mark.gpu
def test_stale_work_lists_are_refused():
    """The flattened lists replay the sub-task recursion with the cells' h_max at the build; the reference
    re-evaluates cell_can_recurse_in_*_hydro_task every step (src/cell_unskip.c:808-815).  When h grows
    past gamma * h_max >= dmin / 2 in a cell the lists recursed through, the step must fail loudly
    instead of silently dropping the neighbours in non-adjacent progeny."""
    from swift_subtask_dev_b200 import SwiftCudaError
    ic = ics.uniform_random(96 ** 3, seed=12)   # 512 particles per top-level cell: split, sub-tasks recurse
    b = HydroBackend(make_params(periodic=True))
    b.build(ic)
    b.step()
    grown = dict(ic, h=(ic["h"] * 2.5).astype(np.float32))
    b.update_particles(grown)
    with pytest.raises(SwiftCudaError, match="work lists are stale"):
        b.step()
    # a rebuild with the new h is the remedy
    b.build(grown)
    b.step()
    b.close()


@pytest.mark.gpu
def test_max_active_bin_can_change_between_steps():
    """Individual time steps through the struct part entries: the work lists hold every cell, which
    particles are active is e->max_active_bin of the step (swift_cuda_set_max_active_bin)."""
    ic = ics.clustered(30000, nclumps=6, seed=14)
    rng = np.random.default_rng(2)
    ic["time_bin"] = rng.integers(1, 5, size=len(ic["h"])).astype(np.int8)
    parts0 = be.parts_from_ic(ic)
    for k in ("rho", "pressure", "soundspeed"):
        parts0[k] = 1.0

    def run(build_bin, step_bin):
        parts = parts0.copy()
        b = HydroBackend(make_params(periodic=True, max_active_bin=build_bin))
        b.build_from_parts(parts)
        if step_bin != build_bin:
            b.set_max_active_bin(step_bin)
        b.step_parts(parts)
        c = b.counters()
        b.close()
        return parts, c
    a, ca = run(2, 2)
    bb, cb = run(56, 2)      # built for everything, stepped with bin 2
    for k in ("h", "rho", "a_hydro", "u_dt", "h_dt", "v_sig", "min_ngb_time_bin"):
        assert np.array_equal(a[k], bb[k]), k
    assert ca["n_force"] == cb["n_force"] and ca["n_density_redo"] == cb["n_density_redo"]
    act = parts0["time_bin"] <= 2
    assert (a["a_hydro"][~act] == 0).all() and (np.abs(a["a_hydro"][act]).sum() > 0)


@pytest.mark.gpu
@pytest.mark.parametrize("fields_mode", [0, 1])
def test_records_in_cell_order_come_back_in_pieces(fields_mode, monkeypatch):
    """s->parts after space_rebuild: every cell's records contiguous.  swift_cuda_step_parts then takes the
    records in pieces (unpack, cell h_max / boxes / sorts and the first density pass of a piece under the copy
    of the next) and runs the force loop over pieces of consecutive leaves, copying each finished piece back
    while the next one runs: the same bits as the one-piece path and as the records in arbitrary order."""
    ic = ics.clustered(300000, nclumps=30, seed=12)
    kw = dict(periodic=True, fields_mode=fields_mode)
    order = be.host_plan(make_params(**kw), ic)["order"]
    ics_sorted = ics.permute(ic, order)
    out = {}
    for name, src, pieces in (("shuffled", ic, "8"), ("sorted_1", ics_sorted, "1"), ("sorted_8", ics_sorted, "8"),
                              ("sorted_5", ics_sorted, "5")):
        monkeypatch.setenv("SWIFT_CUDA_PIECES", pieces)
        parts = be.parts_from_ic(src, pinned=(name == "sorted_8"))
        b = HydroBackend(make_params(**kw))
        b.build_from_parts(parts)
        b.step_parts(parts)
        parts["h"] = src["h"]
        b.step_parts(parts)            # a second step: the pieces' events and the copy stream are reused
        out[name] = (np.array(parts, copy=True), b.counters())
        b.close()
    inv = np.empty(len(order), dtype=np.int64)
    inv[order] = np.arange(len(order))
    base, cbase = out["sorted_1"]
    # the piece path really ran: P force launches and P write-backs instead of one of each
    # the piece path really ran: per piece a density and a force launch, the prep kernels and a write-back
    assert out["sorted_8"][1]["n_kernel_launches"] >= out["sorted_5"][1]["n_kernel_launches"] + 3 * 6
    assert out["sorted_5"][1]["n_kernel_launches"] >= cbase["n_kernel_launches"] + 4 * 6
    assert out["shuffled"][1]["n_kernel_launches"] == cbase["n_kernel_launches"]
    for name in ("sorted_8", "sorted_5", "shuffled"):
        got, cnt = out[name]
        if name == "shuffled":
            got = got[order]
        for k in ("n_density", "n_density_redo", "n_force"):
            assert cnt[k] == cbase[k], (name, k)
        for f in base.dtype.names:
            assert np.array_equal(got[f], base[f], equal_nan=True) if got[f].dtype.kind == "f" else \
                np.array_equal(got[f], base[f]), (name, f)
