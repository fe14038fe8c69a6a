"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every
symbol include/swift_cuda.h declares, its structs match the ctypes mirror, it
refuses to compute without a GPU, and its host-side tree / task builder makes
the same decisions as the oracle's restatement of space_split /
scheduler_splittask_hydro."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib as ol
from swift_subtask_dev_b200 import backend as be
from swift_subtask_dev_b200 import ics, make_params

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "swift_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(swift_cuda_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = be.load_library()
    syms = _declared_symbols()
    assert len(syms) >= 25
    for name in syms:
        assert hasattr(lib, name), "libswiftcuda.so does not export %s" % name
    assert sorted(be.EXPORTS) == syms, "backend.EXPORTS and include/swift_cuda.h disagree"
    assert lib.swift_cuda_abi_version() == 4


def test_struct_layouts_match_ctypes_mirror():
    lib = be.load_library()
    for which, cls in enumerate([be.Params, be.Cell, be.Task, be.PartsIn, be.PartsOut, be.Counters, be.Timings,
                                 be.Til, be.Integrator, be.StateOut, be.PartLayout, be.SphenixProps, be.SphenixIn,
                                 be.SphenixOut]):
        assert lib.swift_cuda_sizeof(which) == C.sizeof(cls), cls.__name__
    # the oracle's tree structs are layout-compatible (tests memmove between them)
    assert C.sizeof(ol.Cell) == C.sizeof(be.Cell) and C.sizeof(ol.Task) == C.sizeof(be.Task)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(be.SwiftCudaError) as e:
        be.HydroBackend(make_params())
    assert "no CUDA device" in str(e.value)


CASES = [
    ("uniform64k", lambda: ics.uniform_random(40 ** 3, seed=3), dict(periodic=True)),
    ("uniform_split", lambda: ics.uniform_random(96 ** 3, seed=5), dict(periodic=True)),
    ("open_box", lambda: ics.uniform_random(30 ** 3, seed=7), dict(periodic=False)),
    ("clustered", lambda: ics.clustered(60000, nclumps=16, seed=3), dict(periodic=True)),
]


@pytest.mark.parametrize("name,make,kw", CASES, ids=[c[0] for c in CASES])
def test_host_plan_matches_oracle_tree_and_tasks(name, make, kw):
    ic = make()
    plan = be.host_plan(make_params(**kw), ic)
    order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(ol.make_params(**kw), ic["x"], ic["h"])
    assert plan["cdim"] == tuple(cdim)
    assert plan["ncells"] == len(cells) and plan["ntasks"] == nt
    # the two builders may number the progeny differently: compare by (first, count, depth)
    def key(c):
        return (c.first, c.count, c.depth)

    def canon(cs, n):
        return sorted((key(cs[k]), tuple(cs[k].loc), tuple(cs[k].width), cs[k].split,
                       key(cs[cs[k].parent]) if cs[k].parent >= 0 else None,
                       tuple(key(cs[p]) if p >= 0 else None for p in cs[k].progeny)) for k in range(n))
    assert canon(plan["cells"], plan["ncells"]) == canon(cells, len(cells)), "cells differ"

    def canon_tasks(ts, n, cs):
        return sorted((ts[k].type, key(cs[ts[k].ci]), key(cs[ts[k].cj]) if ts[k].cj >= 0 else None)
                      for k in range(n))
    assert canon_tasks(plan["tasks"], nt, plan["cells"]) == canon_tasks(tasks, nt, cells), "task lists differ"
    # same particles in every leaf
    for k in range(0, len(cells), max(1, len(cells) // 200)):
        c = cells[k]
        assert set(plan["order"][c.first:c.first + c.count]) == set(order[c.first:c.first + c.count])
    assert plan["n_interactions"] > 0 and plan["n_items"] >= plan["n_interactions"]


def test_host_plan_rejects_too_large_h_in_periodic_box():
    ic = ics.uniform_random(1000, seed=1)
    ic["h"][:] = 0.3  # 1.1 * gamma * h > box / 3
    with pytest.raises(be.SwiftCudaError):
        be.host_plan(make_params(periodic=True), ic)


def test_struct_part_mirror_has_the_reference_layout():
    """The numpy mirror of struct part: 128-byte records, the density / force union overlapping as in
    src/hydro/Minimal/hydro_part.h:133-189 (offsets probe-compiled against the reference headers)."""
    for scheme in (be.SCHEME_MINIMAL, be.SCHEME_PLANETARY):
        dt = be.part_dtype(scheme)
        assert dt.itemsize == 128
        off = {k: dt.fields[k][1] for k in dt.names}
        assert off["x"] == 16 and off["v"] == 40 and off["a_hydro"] == 52 and off["rho"] == 80
        assert off["wcount"] == off["f"] == 84 and off["rot_v"] == off["h_dt"] == 100
        assert off["time_bin"] == (113 if scheme == be.SCHEME_MINIMAL else 117)
    ic = ics.uniform_random(100, seed=1)
    parts = be.parts_from_ic(ic)
    assert np.array_equal(parts["x"], ic["x"]) and np.array_equal(parts["h"], ic["h"])
    assert (parts["time_bin"] == 1).all() and (parts["rho"] == 0).all()


DIGEST_CASES = [
    ("uniform_split", lambda: ics.uniform_random(96 ** 3, seed=5), dict(periodic=True)),
    ("clustered", lambda: ics.clustered(60000, nclumps=16, seed=3), dict(periodic=True)),
    ("open_small", lambda: ics.uniform_random(20 ** 3, seed=7), dict(periodic=False)),
    ("timebins", lambda: dict(ics.uniform_random(48 ** 3, seed=9),
                              time_bin=np.random.default_rng(2).integers(1, 5, 48 ** 3).astype(np.int8)),
     dict(periodic=True, max_active_bin=2)),
]


@pytest.mark.parametrize("name,make,kw", DIGEST_CASES, ids=[c[0] for c in DIGEST_CASES])
def test_host_builder_threads_give_the_same_lists(name, make, kw, monkeypatch):
    """The host builder shares its counting sort, the octree split, the per-leaf interaction lists and the lane
    assignment out over host threads; the tree, the task list, the cell order, both item lists, units, lanes and
    sort segments must come out byte for byte as from the single-threaded form (SWIFT_CUDA_HOST_THREADS=1)."""
    ic = make()
    prm = make_params(**kw)
    monkeypatch.setenv("SWIFT_CUDA_HOST_THREADS", "1")
    serial = be.host_plan_digest(prm, ic)
    for nthreads in ("2", "5", "16"):
        monkeypatch.setenv("SWIFT_CUDA_HOST_THREADS", nthreads)
        assert be.host_plan_digest(prm, ic) == serial, "%s host threads" % nthreads
    monkeypatch.delenv("SWIFT_CUDA_HOST_THREADS")
    assert be.host_plan_digest(prm, ic) == serial


def test_host_builder_takes_degenerate_inputs():
    """No particle, one, two, and sets without a single active particle (what the reference's testActivePair hands to
    the pair tasks): the host builder returns a plan, the same with one host thread and with all."""
    import parity
    from swift_subtask_dev_b200 import backend as be

    def ic_n(n, tb=None):
        rng = np.random.default_rng(1)
        ic = {"x": rng.random((n, 3)) * 3.0, "v": np.zeros((n, 3), np.float32), "mass": np.ones(n, np.float32),
              "h": np.full(n, 0.2, np.float32), "u": np.ones(n, np.float32), "id": np.arange(n, dtype=np.int64)}
        if tb is not None:
            ic["time_bin"] = np.full(n, tb, np.int8)
        return ic

    prm = parity.gpu_params(dict(periodic=False, dim=(3.0, 3.0, 3.0), max_active_bin=56))
    seen = set()
    for n, tb in ((0, None), (1, None), (2, None), (2, 57), (500, 57), (500, None)):
        d = be.host_plan_digest(prm, ic_n(n, tb))
        os.environ["SWIFT_CUDA_HOST_THREADS"] = "1"
        try:
            assert be.host_plan_digest(prm, ic_n(n, tb)) == d
        finally:
            del os.environ["SWIFT_CUDA_HOST_THREADS"]
        seen.add(d)
    assert len(seen) == 6
