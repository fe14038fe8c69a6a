"""Table equations of state on the GPU against the fixtures the REFERENCE's own compiled operators, table loaders and
interpolation wrote (tests/golden/eos_tables/impact3k_tables*.npz, tests/golden/make_golden_eos_step.py):
python scripts/eos_tables_gpu_check.py [step|run|all]

  step  one whole step of a two-body impact whose materials live in five synthetic tables (SESAME-like, ANEOS-like,
        custom, two HM80-like) next to Tillotson iron: pair lists, counts and every field bit for bit
  run   three engine steps with the particles resident on the device (the drift and kick2 evaluate the tables too):
        the final state and the log against the reference's

tests/test_zyz_gpu_eos_tables.py runs it in a process of its own.  Exits 0 only when all checks pass."""
import ast
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import golden_util as gu  # noqa: E402
import integ_lib as il  # noqa: E402
import oracle_lib as ol  # noqa: E402
import parity  # noqa: E402
from swift_subtask_dev_b200 import HydroBackend  # noqa: E402
from swift_subtask_dev_b200 import backend as B  # noqa: E402

DIR = os.path.join(ROOT, "tests", "golden", "eos_tables")
TABLES = {300: "sesame_like_iron.txt", 400: "aneos_like_forsterite.txt", 900: "custom_0.txt", 200: "hm80_like_HHe.txt",
          202: "hm80_like_rock.txt"}
FAILED = []


def check(ok, what):
    print(("ok   " if ok else "FAIL ") + what, flush=True)
    if not ok:
        FAILED.append(what)
    return ok


def backend_with_tables(kw, binary_dir=None):
    units = np.load(os.path.join(DIR, "reference_results.npz"))["units"]
    be = HydroBackend(parity.gpu_params(kw, ol.eos_from_json()))
    for mat_id, fname in TABLES.items():
        t = B.load_eos_table(os.path.join(DIR, fname), mat_id, tuple(units))
        if binary_dir is not None:
            # through the flat binary form, as a production run would read it
            path = os.path.join(binary_dir, "%d.bin" % mat_id)
            assert be.lib.swift_cuda_eos_table_write_bin(path.encode(), B.C.byref(t)) == 0
            B.free_eos_table(t)
            t = B.load_eos_table(path, mat_id, tuple(units))
        be.set_eos_table(t)
        B.free_eos_table(t)     # the context has its own copy on the device
    return be


def load_step_fixture():
    z = np.load(os.path.join(DIR, "impact3k_tables.npz"))
    g = {k: z[k] for k in z.files}
    g["params"] = ast.literal_eval(str(g["params"]))
    g["ic"] = {"x": g["x"], "v": g["v"], "mass": g["mass"], "h": g["h"], "u": g["u"], "id": g["id"],
               "mat_id": g["mat_id"], "time_bin": None}
    return g


def step_checks():
    import tempfile
    g = load_step_fixture()
    cells, tasks, nt = gu.tree_structs(g)
    n = len(g["h"])
    for binary in (False, True):
        with tempfile.TemporaryDirectory() as tmp:
            be = backend_with_tables(g["params"], tmp if binary else None)
        be.build_from_cells(g["ic"], cells, tasks, nt)
        be.record_pairs(n * 400)
        be.step()
        out = be.results(first_pass=True)
        out["counters"] = be.counters()
        out["pairs"] = [be.fetch_pairs(k, n * 400) for k in range(3)]
        be.close()
        ref = {k: g[k] for _, k in parity.FIELD_MAP}
        ref["counts"] = [int(v) for v in g["counts"]]
        r = parity.compare(out, ref, verbose=False)
        tag = "tables read from the flat binary form" if binary else "tables read from the text files"
        for s in range(3):
            check(gu.pair_digest(out["pairs"][s]) == str(g["digests"][s]), "step (%s): pair list of loop %d" % (tag, s))
        try:
            parity.assert_exact(r)
            check(True, "step (%s): counts and every field bit for bit; P, c of %d table particles" % (
                tag, int((g["mat_id"] >= 200).sum())))
        except AssertionError as e:
            check(False, "step (%s): %s" % (tag, str(e)[:300]))
        for m in np.unique(g["mat_id"]):
            sel = g["mat_id"] == m
            bad = int((np.asarray(out["pressure"])[sel] != g["g_pressure"][sel]).sum() +
                      (np.asarray(out["soundspeed"])[sel] != g["g_soundspeed"][sel]).sum())
            check(bad == 0, "step (%s): material %d, %d particles, %d P / c values differ" % (tag, int(m), int(sel.sum()), bad))


def run_checks():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden_eos_step as gen
    icc, tree, kw, ip = gen.case()
    z = np.load(os.path.join(DIR, "impact3k_tables_run.npz"))
    be = backend_with_tables(dict(kw, max_active_bin=il.NUM_TIME_BINS))
    sg, lg = il.run_loop_gpu(be, icc, ip, tree, len(z["log_int"]) - 1)
    be.close()
    bad = [k for k in il.STATE_KEYS if not np.array_equal(np.asarray(sg[k]).reshape(-1), z[k].reshape(-1))]
    check(not bad, "run: final state after %d engine steps against the reference's %s" % (len(z["log_int"]) - 1, bad))
    check([[l[0], l[1], l[2], l[5]] for l in lg] == z["log_int"].tolist(), "run: active counts and sync points")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("step", "all"):
        step_checks()
    if what in ("run", "all"):
        run_checks()
    print("eos_tables_gpu_check:", "all ok" if not FAILED else "%d FAILED" % len(FAILED), flush=True)
    sys.exit(1 if FAILED else 0)
