"""TEST-ONLY bindings for the table equations of state: the reference's own loaders and
interpolation (oracle/_ref/libswiftref_planetary.so: ref_eos_table_*), the library's host-side loaders
(libswiftcuda.so: swift_cuda_eos_table_*, no GPU needed) and the device's interpolation source compiled for the
host (tests/host_kernels/eos_table_host.cpp around swift_subtask_dev_b200/csrc/eos_table.h)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_FP = C.POINTER(C.c_float)


from oracle_lib import EosTable as Table  # noqa: E402  oracle_eos_table == swift_cuda_eos_table, field for field


class Units(C.Structure):
    _fields_ = [("density", C.c_double), ("energy_per_unit_mass", C.c_double), ("pressure", C.c_double),
                ("speed", C.c_double)]


SCALARS = ("kind", "mat_id", "num_rho", "num_T", "P_tiny", "c_tiny", "log_rho_min", "log_rho_step", "inv_log_rho_step",
           "log_u_min", "log_u_step", "inv_log_u_step", "bulk_mod", "P_min_for_c_min")


def table_arrays(t):
    """dict of numpy copies of a Table's arrays and scalars."""
    out = {k: getattr(t, k) for k in SCALARS}
    n2 = t.num_rho * t.num_T
    out["P"] = np.ctypeslib.as_array(t.P, shape=(n2,)).copy()
    if t.kind == 1:
        out["log_rho"] = np.ctypeslib.as_array(t.log_rho, shape=(t.num_rho,)).copy()
        out["log_u"] = np.ctypeslib.as_array(t.log_u, shape=(n2,)).copy()
        out["c"] = np.ctypeslib.as_array(t.c, shape=(n2,)).copy()
    return out


def table_from_arrays(a):
    """A Table over numpy arrays (kept alive in the returned list)."""
    t = Table()
    keep = []
    for k in SCALARS:
        setattr(t, k, int(a[k]) if k in ("kind", "mat_id", "num_rho", "num_T") else float(a[k]))
    for k in ("log_rho", "log_u", "P", "c"):
        if k in a and a[k] is not None and len(np.atleast_1d(a[k])):
            arr = np.ascontiguousarray(a[k], dtype=np.float32)
            keep.append(arr)
            setattr(t, k, arr.ctypes.data_as(_FP))
    return t, keep


def ref_lib():
    import oracle_lib as ol
    return ol.ref("planetary")


def ref_units():
    r = (C.c_double * 4)()
    assert ref_lib().ref_eos_unit_ratios(r) == 0
    return Units(*[float(v) for v in r])


def ref_load(mat_id, path):
    lib = ref_lib()
    rc = lib.ref_eos_table_load(mat_id, path.encode())
    assert rc == 0, rc
    t = Table()
    assert lib.ref_eos_table_get(mat_id, C.byref(t)) == 0
    return t


def ref_eval(mat_id, rho, u):
    rho = np.ascontiguousarray(rho, dtype=np.float32)
    u = np.ascontiguousarray(u, dtype=np.float32)
    P, c = np.zeros_like(rho), np.zeros_like(rho)
    rc = ref_lib().ref_eos_table_eval(mat_id, len(rho), rho.ctypes.data_as(_FP), u.ctypes.data_as(_FP),
                                      P.ctypes.data_as(_FP), c.ctypes.data_as(_FP))
    assert rc == 0
    return P, c


_host = None


def host_body():
    """The device's interpolation source, built for the host."""
    global _host
    if _host is None:
        import tempfile
        so = os.path.join(tempfile.mkdtemp(prefix="eos_host_"), "libeostab.so")
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-o", so,
                        os.path.join(ROOT, "tests", "host_kernels", "eos_table_host.cpp")], check=True)
        _host = C.CDLL(so)
    return _host


def host_eval(t, rho, u):
    rho = np.ascontiguousarray(rho, dtype=np.float32)
    u = np.ascontiguousarray(u, dtype=np.float32)
    P, c = np.zeros_like(rho), np.zeros_like(rho)
    host_body().eos_table_eval_host(C.byref(t), len(rho), rho.ctypes.data_as(_FP), u.ctypes.data_as(_FP),
                                    P.ctypes.data_as(_FP), c.ctypes.data_as(_FP))
    return P, c


def product_lib():
    from swift_subtask_dev_b200 import load_library
    lib = load_library()
    lib.swift_cuda_eos_table_load_sesame.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.swift_cuda_eos_table_load_hm80.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.swift_cuda_eos_table_write_bin.argtypes = [C.c_char_p, C.c_void_p]
    lib.swift_cuda_eos_table_read_bin.argtypes = [C.c_char_p, C.c_void_p]
    lib.swift_cuda_eos_table_free.argtypes = [C.c_void_p]
    lib.swift_cuda_last_error.restype = C.c_char_p
    return lib


def sample_points(a, units_density, seed=1, n=4000):
    """(rho, u) in internal units over the table: inside it in density (outside, the reference reads beyond its
    arrays, sesame.h:499-503), from below to above it in u, plus table nodes themselves and u <= 0."""
    rng = np.random.default_rng(seed)
    if a["kind"] == 1:
        lr0, lr1 = float(a["log_rho"][0]), float(a["log_rho"][-1])
        lu = a["log_u"][np.isfinite(a["log_u"])]
        lu0, lu1 = float(lu.min()), float(lu.max())
    else:
        lr0 = a["log_rho_min"]
        lr1 = lr0 + a["log_rho_step"] * (a["num_rho"] - 1)
        lu0 = a["log_u_min"]
        lu1 = lu0 + a["log_u_step"] * (a["num_T"] - 1)
        lr0, lr1 = lr0 - 1.0, lr1 + 1.0     # the regular grid is extrapolated safely on both sides
    rho = np.exp(rng.uniform(lr0, lr1, n)).astype(np.float32)
    u = np.exp(rng.uniform(lu0 - 2.0, lu1 + 2.0, n)).astype(np.float32)
    if a["kind"] == 1:
        rho = np.clip(rho, np.exp(np.float32(lr0)) * np.float32(1.0001), np.exp(np.float32(lr1)) * np.float32(0.9999))
        # exact table nodes
        k = rng.integers(0, a["num_rho"] - 1, 200)
        j = rng.integers(0, a["num_T"], 200)
        rho[:200] = np.exp(a["log_rho"][k])
        u[:200] = np.exp(a["log_u"].reshape(a["num_rho"], a["num_T"])[k, j])
        rho[:200] = np.clip(rho[:200], np.exp(np.float32(lr0)) * np.float32(1.0001), None)
    u[200:220] = 0.0
    u[220:230] = -1.0
    return rho.astype(np.float32), u.astype(np.float32)


def eos_with_tables(eos, tables):
    """An oracle_lib.Eos (ideal gas + Tillotson, e.g. ol.eos_from_json()) that also knows the given prepared tables
    (dicts as table_arrays / the fixture's "<name>/table/*" entries).  Returns (eos, keep-alive list)."""
    keep = []
    arr = (Table * max(len(tables), 1))()
    for k, a in enumerate(tables):
        t, kp = table_from_arrays(a)
        keep.append(kp)
        C.memmove(C.byref(arr, k * C.sizeof(Table)), C.byref(t), C.sizeof(Table))
    keep.append(arr)
    eos.n_tab = len(tables)
    eos.tab = C.cast(arr, C.POINTER(Table))
    return eos, keep


def fixture_tables(gold=None):
    """The prepared tables of tests/golden/eos_tables/reference_results.npz, one dict per table."""
    if gold is None:
        z = np.load(os.path.join(ROOT, "tests", "golden", "eos_tables", "reference_results.npz"))
        gold = {k: z[k] for k in z.files}
    names = sorted({k.split("/")[0] for k in gold if "/table/" in k})
    out = []
    for name in names:
        pre = name + "/table/"
        out.append({k[len(pre):]: gold[k] for k in gold if k.startswith(pre)})
    return out
