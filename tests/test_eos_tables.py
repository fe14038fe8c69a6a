"""Table equations of state of the Planetary scheme, CPU side.  The reference's table FILES are
not in its tree, so parity is pinned on synthetic tables in the reference's file formats
(tests/golden/eos_tables/*.txt, tests/golden/make_eos_tables.py) that the REFERENCE's own code read and evaluated
(load_table_* / prepare_table_* / convert_units_* and SESAME_* / HM80_*_from_internal_energy behind
oracle/ref_harness.c; results in tests/golden/eos_tables/reference_results.npz):

  * the library's host-side loaders (swift_cuda_eos_table_load_sesame / _hm80, no GPU needed) give the
    reference's prepared arrays bit for bit, and survive a round trip through the flat binary form;
  * the device's interpolation source (swift_subtask_dev_b200/csrc/eos_table.h, the file nvcc compiles into
    eos_eval), built for the host by g++ without contraction, gives the reference's P and c bit for bit on 4000
    points per table -- inside, below and above the table in u, on table nodes, at u <= 0;
  * its logf / expf agree with this machine's libm (what the reference calls);
  * live against the compiled reference where oracle/_ref is present;
  * the C oracle (oracle/sph_oracle.c, its own restatement of the table functions) reproduces the whole step and
    the three-step run the compiled reference made with table materials (impact3k_tables*.npz).

GPU tests of the same tables: tests/test_zyz_gpu_eos_tables.py."""
import ctypes as C
import os

import numpy as np
import pytest

import eos_lib as el
import oracle_lib as ol

DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "eos_tables")
TABLES = {
    "sesame_like_iron": ("sesame_like_iron.txt", "sesame", 300, 20220714),
    "aneos_like_forsterite": ("aneos_like_forsterite.txt", "sesame", 400, 20220714),
    "custom_0": ("custom_0.txt", "sesame", 900, 0),
    "hm80_like_HHe": ("hm80_like_HHe.txt", "hm80", 200, 20201003),
    "hm80_like_rock": ("hm80_like_rock.txt", "hm80", 202, 20201003),
}


@pytest.fixture(scope="module")
def gold():
    z = np.load(os.path.join(DIR, "reference_results.npz"))
    return {k: z[k] for k in z.files}


def gold_table(gold, name):
    pre = name + "/table/"
    return {k[len(pre):]: gold[k] for k in gold if k.startswith(pre)}


def load_product(name, units):
    fname, fmt, mat_id, date = TABLES[name]
    lib = el.product_lib()
    t = el.Table()
    path = os.path.join(DIR, fname).encode()
    if fmt == "sesame":
        rc = lib.swift_cuda_eos_table_load_sesame(path, mat_id, date, C.byref(units), C.byref(t))
    else:
        rc = lib.swift_cuda_eos_table_load_hm80(path, mat_id, C.byref(units), C.byref(t))
    assert rc == 0, lib.swift_cuda_last_error().decode()
    return lib, t


def assert_tables_equal(a, b, what):
    for k in el.SCALARS:
        assert np.float32(a[k]) == np.float32(b[k]) or (np.isnan(a[k]) and np.isnan(b[k])), "%s: %s %r != %r" % (
            what, k, a[k], b[k])
    for k in ("log_rho", "log_u", "P", "c"):
        if k in a or k in b:
            x, y = np.asarray(a[k], dtype=np.float32), np.asarray(b[k], dtype=np.float32)
            assert x.shape == y.shape and np.array_equal(x.view(np.uint32), y.view(np.uint32)), "%s: array %s" % (what, k)


@pytest.mark.parametrize("name", list(TABLES))
def test_loader_reproduces_the_references_prepared_table(name, gold, tmp_path):
    units = el.Units(*[float(v) for v in gold["units"]])
    lib, t = load_product(name, units)
    a = el.table_arrays(t)
    assert_tables_equal(a, gold_table(gold, name), name)
    # flat binary form: there and back
    path = str(tmp_path / (name + ".bin")).encode()
    assert lib.swift_cuda_eos_table_write_bin(path, C.byref(t)) == 0
    t2 = el.Table()
    assert lib.swift_cuda_eos_table_read_bin(path, C.byref(t2)) == 0, lib.swift_cuda_last_error().decode()
    assert_tables_equal(el.table_arrays(t2), a, name + " (binary round trip)")
    n2 = a["num_rho"] * a["num_T"]
    assert os.path.getsize(path) == 64 + 4 * (n2 if a["kind"] == 2 else a["num_rho"] + 3 * n2)
    # a truncated file is refused, not half read
    with open(path, "rb") as f:
        blob = f.read()
    with open(path, "wb") as f:
        f.write(blob[:-8])
    t3 = el.Table()
    assert lib.swift_cuda_eos_table_read_bin(path, C.byref(t3)) != 0
    assert b"truncated" in lib.swift_cuda_last_error()
    lib.swift_cuda_eos_table_free(C.byref(t))
    lib.swift_cuda_eos_table_free(C.byref(t2))
    assert not t.P and t.kind == 0


def test_loader_refusals(gold, tmp_path):
    units = el.Units(*[float(v) for v in gold["units"]])
    lib = el.product_lib()
    t = el.Table()
    p = os.path.join(DIR, "sesame_like_iron.txt").encode()
    assert lib.swift_cuda_eos_table_load_sesame(p, 300, 20230101, C.byref(units), C.byref(t)) != 0
    assert b"version_date" in lib.swift_cuda_last_error()           # the reference's error(), sesame.h:163-169
    assert lib.swift_cuda_eos_table_load_sesame(b"/nonexistent/table.txt", 300, 0, C.byref(units), C.byref(t)) != 0
    assert b"Failed to open" in lib.swift_cuda_last_error()
    assert lib.swift_cuda_eos_table_load_hm80(os.path.join(DIR, "hm80_like_HHe.txt").encode(), 300, C.byref(units),
                                              C.byref(t)) != 0
    bad = tmp_path / "short.txt"
    bad.write_text(open(os.path.join(DIR, "custom_0.txt")).read()[:2000])
    assert lib.swift_cuda_eos_table_load_sesame(str(bad).encode(), 900, 0, C.byref(units), C.byref(t)) != 0
    assert b"Failed to read" in lib.swift_cuda_last_error()
    assert not t.P and not t.log_u                                   # nothing half-allocated is handed out
    assert lib.swift_cuda_eos_table_read_bin(p, C.byref(t)) != 0     # a text table is not a binary one


@pytest.mark.parametrize("name", list(TABLES))
def test_device_interpolation_source_on_the_host_reproduces_the_reference(name, gold):
    t, keep = el.table_from_arrays(gold_table(gold, name))
    P, c = el.host_eval(t, gold[name + "/rho"], gold[name + "/u"])
    wantP, wantc = gold[name + "/P"], gold[name + "/c"]
    badP = P.view(np.uint32) != wantP.view(np.uint32)
    badc = c.view(np.uint32) != wantc.view(np.uint32)
    assert not badP.any(), "%d pressures differ, first at rho %r u %r: %r != %r" % (
        badP.sum(), gold[name + "/rho"][badP][0], gold[name + "/u"][badP][0], P[badP][0], wantP[badP][0])
    assert not badc.any(), "%d sound speeds differ" % badc.sum()
    assert (P > 0).sum() > len(P) // 2 and (c > 0).sum() > len(c) // 2
    assert (P[gold[name + "/u"] <= 0] == 0).all()


def test_logf_expf_as_libm():
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-100, 90, 200000), np.exp(rng.uniform(-100, 88, 400000)),
                        [0.0, 1.0, np.inf, 1e-45, 1.17549435e-38, 3.4028235e38, -1.0, np.nan]]).astype(np.float32)
    lo, ex = np.zeros_like(x), np.zeros_like(x)
    el.host_body().eos_table_logf_expf_host(len(x), x.ctypes.data_as(el._FP), lo.ctypes.data_as(el._FP),
                                            ex.ctypes.data_as(el._FP))
    libm = C.CDLL("libm.so.6")
    libm.logf.restype, libm.logf.argtypes = C.c_float, [C.c_float]
    libm.expf.restype, libm.expf.argtypes = C.c_float, [C.c_float]
    sample = np.concatenate([np.arange(0, len(x), 37), np.arange(len(x) - 8, len(x))])
    for k in sample:
        a, b = np.float32(libm.logf(float(x[k]))), lo[k]
        assert a.view(np.uint32) == b.view(np.uint32) or (np.isnan(a) and np.isnan(b)), ("logf", x[k], a, b)
        a, b = np.float32(libm.expf(float(x[k]))), ex[k]
        assert a.view(np.uint32) == b.view(np.uint32) or (np.isnan(a) and np.isnan(b)), ("expf", x[k], a, b)


@pytest.mark.skipif(not ol.ref_available("planetary"), reason="oracle/_ref not built (needs the reference tree)")
def test_live_against_the_compiled_reference(gold):
    """Other points than the fixture holds, through the whole chain: the library's loader, then the device's
    interpolation source on the host, against the reference's loader and functions."""
    units = el.ref_units()
    assert [units.density, units.energy_per_unit_mass, units.pressure, units.speed] == [float(v) for v in gold["units"]]
    for name, (fname, fmt, mat_id, date) in TABLES.items():
        el.ref_load(mat_id, os.path.join(DIR, fname))
        lib, t = load_product(name, units)
        rho, u = el.sample_points(el.table_arrays(t), units.density, seed=99, n=20000)
        P, c = el.host_eval(t, rho, u)
        rP, rc = el.ref_eval(mat_id, rho, u)
        assert np.array_equal(P.view(np.uint32), rP.view(np.uint32)), name
        assert np.array_equal(c.view(np.uint32), rc.view(np.uint32)), name
        lib.swift_cuda_eos_table_free(C.byref(t))


# ---- the C oracle (oracle/sph_oracle.c: tab_sesame, tab_hm80_pressure, tab_eval) ------------------------------
def _oracle_eos(gold):
    eos, keep = el.eos_with_tables(ol.eos_from_json(), el.fixture_tables(gold))
    return eos, keep


def test_oracle_step_with_table_materials_reproduces_reference_golden(gold):
    """One whole step of the two-body impact whose materials live in the five tables (+ Tillotson iron), by the
    oracle with its own restatement of the table functions, against what the compiled reference wrote
    (tests/golden/make_golden_eos_step.py): pair lists, counts, every field."""
    import ast
    import golden_util as gu
    import parity
    z = np.load(os.path.join(DIR, "impact3k_tables.npz"))
    g = {k: z[k] for k in z.files}
    kw = ast.literal_eval(str(g["params"]))
    ic = {"x": g["x"], "v": g["v"], "mass": g["mass"], "h": g["h"], "u": g["u"], "id": g["id"], "mat_id": g["mat_id"]}
    cells, tasks, nt = gu.tree_structs(g)
    eos, keep = _oracle_eos(gold)
    prm = ol.make_params(record=True, **kw)
    r = ol.run_step("oracle", prm, ic, cells, tasks, nt, record=True, eos=eos)
    assert r["counts"] == [int(v) for v in g["counts"]]
    for s in range(3):
        assert gu.pair_digest(r["pairs"][s]) == str(g["digests"][s])
    for k in list(ol.OUT_FIELDS) + ["f_min_ngb_time_bin"]:
        a, b = np.asarray(r[k]), g[k]
        assert np.array_equal(a, b) or np.array_equal(a.view(np.uint32), b.view(np.uint32)), k
    assert (g["mat_id"] >= 200).sum() > 2900


def test_oracle_run_with_table_materials_reproduces_reference_golden(gold):
    """Three engine steps: the drift (hydro_predict_extra) and kick2 (hydro_reset_predicted_values) evaluate the
    tables too."""
    import sys
    sys.path.insert(0, os.path.dirname(DIR))
    import make_golden_eos_step as gen
    icc, tree, kw, ip = gen.case()
    z = np.load(os.path.join(DIR, "impact3k_tables_run.npz"))
    eos, keep = _oracle_eos(gold)
    s, log = il_run(icc, kw, ip, tree, len(z["log_int"]) - 1, eos)
    import integ_lib as il
    for k in il.STATE_KEYS:
        assert np.array_equal(np.asarray(s[k]).reshape(-1), z[k].reshape(-1)), k
    assert [[l[0], l[1], l[2], l[5]] for l in log] == z["log_int"].tolist()


def il_run(icc, kw, ip, tree, nsteps, eos):
    import integ_lib as il
    return il.run_loop("oracle", icc, kw, ip, tree, nsteps, eos=eos, variant="planetary")
