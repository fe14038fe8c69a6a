"""Python host-side mirror of the C ABI in include/swift_cuda.h (ctypes).

The product is the shared library `libswiftcuda.so` (hand-written sm_100a
kernels + C++ host tree/work-list builder).  This module only marshals numpy
arrays into it; the operator names follow the reference's runner surface
(src/runner_doiact_hydro.h:168-191, src/runner.h:94-119).  There is no CPU
fallback: a missing library or GPU raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "libswiftcuda.so")
CSRC = os.path.join(_HERE, "csrc")

SCHEME_MINIMAL, SCHEME_PLANETARY, SCHEME_SPHENIX = 0, 1, 2
FIELDS_EXACT, FIELDS_FAST = 0, 1

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared", "-fmad=false"]


def build_library(verbose=False, force=False):
    """Compiles the CUDA backend in-tree with nvcc for sm_100a."""
    srcs = [os.path.join(CSRC, f) for f in ("swift_cuda.cu", "host_tree.cpp", "eos_loader.cpp", "particle_io.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("kernels.cuh", "leaf_kernels.cuh", "list_kernels.cuh", "force_fast.cuh", "integrate.cuh", "pack.cuh", "sphenix.cuh", "device_types.cuh", "limit_part.h", "eos_table.h", "host_tree.hpp")] + \
        [os.path.join(ROOT, "include", "swift_cuda.h")]
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    extra = os.environ.get("SWIFT_CUDA_NVCC_EXTRA", "").split()  # tuning experiments (-DSWIFT_MINB_DENSITY=5 ...)
    cmd = ["nvcc"] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB_PATH


class Til(C.Structure):
    _fields_ = [(k, C.c_float) for k in
                ("rho_0", "a", "b", "A", "B", "u_0", "u_iv", "u_cv", "alpha", "beta", "eta_min", "eta_zero", "P_min")] \
        + [("mat_id", C.c_int)]


class Params(C.Structure):
    _fields_ = [("dim", C.c_double * 3), ("periodic", C.c_int), ("eta_neighbours", C.c_float),
                ("h_tolerance", C.c_float), ("h_max", C.c_float), ("h_min", C.c_float),
                ("max_smoothing_iterations", C.c_int), ("viscosity_alpha", C.c_float),
                ("max_active_bin", C.c_int), ("scheme", C.c_int), ("cosmo_a", C.c_float), ("cosmo_H", C.c_float),
                ("n_til", C.c_int), ("til", Til * 8), ("idg_gamma", C.c_float * 4), ("fields_mode", C.c_int)]


class Cell(C.Structure):
    _fields_ = [("loc", C.c_double * 3), ("width", C.c_double * 3), ("first", C.c_int), ("count", C.c_int),
                ("parent", C.c_int), ("progeny", C.c_int * 8), ("depth", C.c_int), ("split", C.c_int)]


class Task(C.Structure):
    _fields_ = [("type", C.c_int), ("ci", C.c_int), ("cj", C.c_int)]


_FP = C.POINTER(C.c_float)


class PartsIn(C.Structure):
    _fields_ = [("n", C.c_int64), ("x", C.POINTER(C.c_double)), ("v", _FP), ("mass", _FP), ("h", _FP), ("u", _FP),
                ("mat_id", C.POINTER(C.c_int32)), ("time_bin", C.POINTER(C.c_int8)), ("id", C.POINTER(C.c_int64)),
                ("rho", _FP), ("pressure", _FP), ("f", _FP), ("soundspeed", _FP), ("balsara", _FP)]


OUT_FIELDS = [("h", 1), ("rho", 1), ("pressure", 1), ("soundspeed", 1), ("f", 1), ("balsara", 1), ("a", 3),
              ("u_dt", 1), ("h_dt", 1), ("v_sig", 1)]
OUT_FIRST = [("d_rho", 1), ("d_rho_dh", 1), ("d_wcount", 1), ("d_wcount_dh", 1), ("d_div_v", 1), ("d_rot_v", 3)]


class PartsOut(C.Structure):
    # field order must match swift_cuda_parts_out
    _fields_ = [("h", _FP), ("rho", _FP), ("pressure", _FP), ("soundspeed", _FP), ("f", _FP), ("balsara", _FP),
                ("a", _FP), ("u_dt", _FP), ("h_dt", _FP), ("v_sig", _FP), ("min_ngb_time_bin", C.POINTER(C.c_int8)),
                ("d_rho", _FP), ("d_rho_dh", _FP), ("d_wcount", _FP), ("d_wcount_dh", _FP), ("d_div_v", _FP),
                ("d_rot_v", _FP)]


class Counters(C.Structure):
    _fields_ = [("n_density", C.c_longlong), ("n_density_redo", C.c_longlong), ("n_force", C.c_longlong),
                ("ghost_sweeps", C.c_int), ("ghost_unconverged", C.c_int), ("n_leaf_interactions", C.c_longlong),
                ("n_sort_entries", C.c_longlong), ("n_kernel_launches", C.c_int)]


class Timings(C.Structure):
    _fields_ = [("sort_ms", C.c_float), ("density_ms", C.c_float), ("ghost_ms", C.c_float), ("force_ms", C.c_float),
                ("total_ms", C.c_float)]


class Integrator(C.Structure):
    # field order must match swift_cuda_integrator
    _fields_ = [("cfl_condition", C.c_float), ("log_max_h_change", C.c_float), ("minimal_internal_energy", C.c_float),
                ("dt_min", C.c_double), ("dt_max", C.c_double), ("time_base", C.c_double)]


class TimestepStaticArgs(C.Structure):
    # field order must match swift_cuda_timestep_static_args
    _fields_ = [("count", C.c_int64), ("h", C.c_void_p), ("v_sig", C.c_void_p), ("h_dt", C.c_void_p),
                ("min_ngb_time_bin", C.c_void_p), ("time_bin", C.c_void_p), ("ti_current", C.c_longlong),
                ("max_active_bin", C.c_int), ("integ", Integrator)]


TIMESTEP_STATIC_MAX = 16777216   # SWIFT_CUDA_TIMESTEP_STATIC_MAX


STATE_FIELDS = [("x", 3), ("v", 3), ("v_full", 3), ("u", 1), ("u_full", 1), ("h", 1), ("rho", 1), ("pressure", 1),
                ("soundspeed", 1), ("v_sig", 1), ("a", 3), ("u_dt", 1), ("h_dt", 1), ("x_diff", 3), ("x_diff_sort", 3)]


class StateOut(C.Structure):
    # field order must match swift_cuda_state_out
    _fields_ = [("x", C.POINTER(C.c_double))] + [(k, _FP) for k, _ in STATE_FIELDS[1:]] + \
               [("time_bin", C.POINTER(C.c_int8)), ("min_ngb_time_bin", C.POINTER(C.c_int8))]


EXPORTS = ["swift_cuda_abi_version", "swift_cuda_sizeof", "swift_cuda_last_error", "swift_cuda_create", "swift_cuda_destroy",
           "swift_cuda_build_from_particles", "swift_cuda_build_from_cells", "swift_cuda_update_particles",
           "swift_cuda_restore_particles", "swift_cuda_set_transfer_count",
           "swift_cuda_step", "swift_cuda_step_io", "swift_cuda_do_sort", "swift_cuda_do_density", "swift_cuda_do_ghost",
           "swift_cuda_do_force", "swift_cuda_get_results", "swift_cuda_get_counters", "swift_cuda_get_timings",
           "swift_cuda_measure_fp32_peak",
           "swift_cuda_get_tree_size", "swift_cuda_get_tree", "swift_cuda_record_pairs", "swift_cuda_fetch_pairs",
           "swift_cuda_halo_pack", "swift_cuda_halo_unpack", "swift_cuda_halo_bytes_per_particle",
           "swift_cuda_get_touched", "swift_cuda_refresh_cell_hmax",
           "swift_cuda_host_plan", "swift_cuda_host_free", "swift_cuda_host_plan_digest",
           "swift_cuda_integrator_init", "swift_cuda_set_time", "swift_cuda_drift", "swift_cuda_kick",
           "swift_cuda_timestep", "swift_cuda_reset_offsets", "swift_cuda_get_state",
           "swift_cuda_get_integrator_timings",
           "swift_cuda_build_from_cells_parts", "swift_cuda_build_from_parts", "swift_cuda_step_parts",
           "swift_cuda_set_max_active_bin", "swift_cuda_timestep_static",
           "swift_cuda_sphenix_set_props", "swift_cuda_sphenix_set_state", "swift_cuda_sphenix_get_results",
           "swift_cuda_sphenix_get_counter", "swift_cuda_do_gradient", "swift_cuda_do_extra_ghost",
           "swift_cuda_do_limiter_loop", "swift_cuda_do_limiter", "swift_cuda_get_limiter_timings",
           "swift_cuda_eos_table_load_sesame", "swift_cuda_eos_table_load_hm80", "swift_cuda_eos_table_write_bin",
           "swift_cuda_eos_table_read_bin", "swift_cuda_eos_table_free", "swift_cuda_set_eos_table",
           "swift_cuda_particles_write_bin", "swift_cuda_particles_read_bin", "swift_cuda_particles_free"]


class EosTable(C.Structure):
    # field order must match swift_cuda_eos_table
    _fields_ = [("kind", C.c_int), ("mat_id", C.c_int), ("num_rho", C.c_int), ("num_T", C.c_int),
                ("log_rho", C.POINTER(C.c_float)), ("log_u", C.POINTER(C.c_float)), ("P", C.POINTER(C.c_float)),
                ("c", C.POINTER(C.c_float)), ("P_tiny", C.c_float), ("c_tiny", C.c_float),
                ("log_rho_min", C.c_float), ("log_rho_step", C.c_float), ("inv_log_rho_step", C.c_float),
                ("log_u_min", C.c_float), ("log_u_step", C.c_float), ("inv_log_u_step", C.c_float),
                ("bulk_mod", C.c_float), ("P_min_for_c_min", C.c_float)]


class EosUnits(C.Structure):
    # field order must match swift_cuda_eos_units
    _fields_ = [("density", C.c_double), ("energy_per_unit_mass", C.c_double), ("pressure", C.c_double),
                ("speed", C.c_double)]


def load_eos_table(path, mat_id, units, version_date=None):
    """Host only.  Reads a table file in the reference's text format (HM80 for mat_id 200-202, SESAME-style
    otherwise) or, for a path ending in .bin, the library's flat binary form.  units: EosUnits or a 4-tuple
    (density, energy per unit mass, pressure, speed: SI value / internal value).  Returns an EosTable whose arrays
    belong to the library (free_eos_table)."""
    lib = load_library()
    t = EosTable()
    if not isinstance(units, EosUnits):
        units = EosUnits(*[float(v) for v in units])
    if path.endswith(".bin"):
        rc = lib.swift_cuda_eos_table_read_bin(path.encode(), C.byref(t))
    elif 200 <= mat_id < 300:
        rc = lib.swift_cuda_eos_table_load_hm80(path.encode(), mat_id, C.byref(units), C.byref(t))
    else:
        if version_date is None:
            version_date = 0 if mat_id >= 900 else 20220714   # set_custom / set_SESAME_*, set_ANEOS_* (sesame.h:56-106)
        rc = lib.swift_cuda_eos_table_load_sesame(path.encode(), mat_id, version_date, C.byref(units), C.byref(t))
    if rc != 0:
        raise SwiftCudaError(lib.swift_cuda_last_error().decode())
    return t


def free_eos_table(t):
    load_library().swift_cuda_eos_table_free(C.byref(t))


class ParticleFile(C.Structure):
    # field order must match swift_cuda_particle_file
    _fields_ = [("n", C.c_int64), ("dim", C.c_double * 3), ("time", C.c_double), ("fields", C.c_int),
                ("coordinates", C.POINTER(C.c_double)), ("velocities", C.POINTER(C.c_float)),
                ("masses", C.POINTER(C.c_float)), ("smoothing_lengths", C.POINTER(C.c_float)),
                ("internal_energies", C.POINTER(C.c_float)), ("particle_ids", C.POINTER(C.c_longlong)),
                ("material_ids", C.POINTER(C.c_int)), ("densities", C.POINTER(C.c_float)),
                ("accelerations", C.POINTER(C.c_float)), ("pressures", C.POINTER(C.c_float))]


# (member of swift_cuda_particle_file, SWIFT_CUDA_FIELD_* bit, key of the particle dicts used here, dtype, width)
PARTICLE_FIELDS = [("coordinates", 1, "x", np.float64, 3), ("velocities", 2, "v", np.float32, 3),
                   ("masses", 4, "mass", np.float32, 1), ("smoothing_lengths", 8, "h", np.float32, 1),
                   ("internal_energies", 16, "u", np.float32, 1), ("particle_ids", 32, "id", np.int64, 1),
                   ("material_ids", 64, "mat_id", np.int32, 1), ("densities", 128, "rho", np.float32, 1),
                   ("accelerations", 256, "a", np.float32, 3), ("pressures", 512, "pressure", np.float32, 1)]


def write_particles(path, ic, dim, time=0.0):
    """Host only.  Writes the particle dict ic (x, v, mass, h, u, id; optional mat_id, rho, a, pressure) as a flat
    binary file with the field set of the reference's HDF5 files (src/hydro/Planetary/hydro_io.h)."""
    lib = load_library()
    f = ParticleFile()
    f.n = len(ic["h"])
    f.dim[:] = [float(v) for v in dim]
    f.time = float(time)
    keep = []
    for member, bit, key, dt, w in PARTICLE_FIELDS:
        if ic.get(key) is None:
            continue
        a = np.ascontiguousarray(ic[key], dtype=dt)
        assert a.size == f.n * w, key
        keep.append(a)
        f.fields |= bit
        setattr(f, member, a.ctypes.data_as(dict(ParticleFile._fields_)[member]))
    if lib.swift_cuda_particles_write_bin(path.encode(), C.byref(f)) != 0:
        raise SwiftCudaError(lib.swift_cuda_last_error().decode())


def read_particles(path):
    """Host only.  Returns (particle dict, dim, time) of a file written by write_particles / the C entry."""
    lib = load_library()
    f = ParticleFile()
    if lib.swift_cuda_particles_read_bin(path.encode(), C.byref(f)) != 0:
        raise SwiftCudaError(lib.swift_cuda_last_error().decode())
    ic = {"time_bin": None}
    n = int(f.n)
    for member, bit, key, dt, w in PARTICLE_FIELDS:
        if not (f.fields & bit):
            ic[key] = None
            continue
        a = np.ctypeslib.as_array(getattr(f, member), shape=(n * w,)).astype(dt, copy=True) if n else np.zeros(0, dt)
        ic[key] = a.reshape(n, 3) if w == 3 else a
    out = (ic, tuple(f.dim), float(f.time))
    lib.swift_cuda_particles_free(C.byref(f))
    return out


class SphenixProps(C.Structure):
    # field order must match swift_cuda_sphenix_props
    _fields_ = [("viscosity_alpha_max", C.c_float), ("viscosity_alpha_min", C.c_float), ("viscosity_length", C.c_float),
                ("diffusion_alpha", C.c_float), ("diffusion_beta", C.c_float), ("diffusion_alpha_max", C.c_float),
                ("diffusion_alpha_min", C.c_float), ("time_base", C.c_double)]


class SphenixIn(C.Structure):
    _fields_ = [("alpha_visc", _FP), ("alpha_diff", _FP), ("div_v_previous_step", _FP)]


SPHENIX_OUT_FIELDS = ("div_v", "laplace_u", "alpha_visc_max_ngb", "alpha_visc", "alpha_diff", "div_v_previous_step",
                      "div_v_dt")


class SphenixOut(C.Structure):
    _fields_ = [(k, _FP) for k in SPHENIX_OUT_FIELDS]


class PartLayout(C.Structure):
    # field order must match swift_cuda_part_layout
    _fields_ = [(k, C.c_int) for k in
                ("stride", "x", "v", "a_hydro", "mass", "h", "u", "u_dt", "rho", "rho_dh", "wcount", "wcount_dh", "div_v",
                 "rot_v", "f", "pressure", "soundspeed", "v_sig", "h_dt", "balsara", "time_bin", "min_ngb_time_bin", "id",
                 "mat_id", "alpha_visc", "div_v_previous_step", "div_v_dt", "alpha_diff", "laplace_u",
                 "alpha_visc_max_ngb")]


# Byte offsets of the reference's `struct part` (src/hydro/Minimal/hydro_part.h:98-189, Planetary :100-192;
# 128 bytes, all subgrid models NONE), probe-compiled against the reference headers with offsetof(): the C
# shim (swift_subtask_dev_b200/shim/swift_cuda_shim.c) computes the same numbers at compile time.
_PART_COMMON = dict(stride=128, x=16, v=40, a_hydro=52, mass=64, h=68, u=72, u_dt=76, rho=80, rho_dh=92, wcount=84,
                    wcount_dh=88, div_v=96, rot_v=100, f=84, pressure=88, soundspeed=92, v_sig=96, h_dt=100, balsara=104,
                    id=0)
_NO_SPHENIX = dict(alpha_visc=-1, div_v_previous_step=-1, div_v_dt=-1, alpha_diff=-1, laplace_u=-1,
                   alpha_visc_max_ngb=-1)
# SPHENIX (src/hydro/SPHENIX/hydro_part.h, 160 bytes, probe-compiled the same way): p->viscosity {div_v, div_v_dt,
# div_v_previous_step, alpha, v_sig} at 84, p->diffusion {laplace_u, alpha} at 104, the density / force union at 112
_PART_SPHENIX = dict(stride=160, id=0, x=16, v=40, a_hydro=52, mass=64, h=68, u=72, u_dt=76, rho=80, div_v=84,
                     div_v_dt=88, div_v_previous_step=92, alpha_visc=96, v_sig=100, laplace_u=104, alpha_diff=108,
                     wcount=112, wcount_dh=116, rho_dh=120, rot_v=124, f=112, pressure=116, soundspeed=120, h_dt=124,
                     balsara=128, alpha_visc_max_ngb=132, time_bin=137, min_ngb_time_bin=139, mat_id=-1)
PART_LAYOUTS = {SCHEME_MINIMAL: dict(_PART_COMMON, time_bin=113, min_ngb_time_bin=115, mat_id=-1, **_NO_SPHENIX),
                SCHEME_PLANETARY: dict(_PART_COMMON, time_bin=117, min_ngb_time_bin=119, mat_id=112, **_NO_SPHENIX),
                SCHEME_SPHENIX: _PART_SPHENIX}
PHASE_SORT, PHASE_DENSITY, PHASE_GHOST, PHASE_FORCE, PHASE_END_FORCE, PHASE_KEEP = 1, 2, 4, 8, 16, 32
PHASE_ALL = 31
PHASE_GRADIENT, PHASE_EXTRA_GHOST = 64, 128   # SPHENIX only
PHASE_ALL_SPHENIX = PHASE_ALL | PHASE_GRADIENT | PHASE_EXTRA_GHOST


def part_dtype(scheme=SCHEME_MINIMAL):
    """numpy record dtype with the reference's struct part layout (the density / force union as two views)."""
    L = PART_LAYOUTS[scheme]
    names = ["id", "x", "v", "a_hydro", "mass", "h", "u", "u_dt", "rho", "wcount", "wcount_dh", "rho_dh", "div_v", "rot_v",
             "f", "pressure", "soundspeed", "v_sig", "h_dt", "balsara", "time_bin", "min_ngb_time_bin"]
    fmts = {"id": "<i8", "x": ("<f8", 3), "v": ("<f4", 3), "a_hydro": ("<f4", 3), "rot_v": ("<f4", 3), "time_bin": "i1",
            "min_ngb_time_bin": "i1"}
    if L["mat_id"] >= 0:
        names.append("mat_id")
        fmts["mat_id"] = "<i4"
    if L["alpha_visc"] >= 0:
        names += ["alpha_visc", "div_v_previous_step", "div_v_dt", "alpha_diff", "laplace_u", "alpha_visc_max_ngb"]
    return np.dtype({"names": names, "formats": [fmts.get(k, "<f4") for k in names],
                     "offsets": [L[k] for k in names], "itemsize": L["stride"]})


def parts_from_ic(ic, scheme=SCHEME_MINIMAL, pinned=False):
    """An array of struct part records from an initial-conditions dict (sums zero: hydro_init_part)."""
    n = len(ic["h"])
    dt = part_dtype(scheme)
    if pinned:
        import torch
        t = torch.zeros(n * dt.itemsize, dtype=torch.uint8, pin_memory=True)
        parts = t.numpy().view(dt)
        _PINNED_KEEP.append(t)
    else:
        parts = np.zeros(n, dtype=dt)
    parts["id"] = ic["id"] if ic.get("id") is not None else np.arange(n)
    parts["x"] = ic["x"]
    parts["v"] = ic["v"]
    parts["mass"] = ic["mass"]
    parts["h"] = ic["h"]
    parts["u"] = ic["u"]
    parts["time_bin"] = ic["time_bin"] if ic.get("time_bin") is not None else 1
    if "mat_id" in dt.names:
        parts["mat_id"] = ic["mat_id"] if ic.get("mat_id") is not None else 0
    if "alpha_visc" in dt.names:
        # p->viscosity.alpha, p->diffusion.alpha, p->viscosity.div_v_previous_step as a previous step left them
        for k_rec, k_ic in (("alpha_visc", "in_alpha_visc"), ("alpha_diff", "in_alpha_diff"),
                            ("div_v_previous_step", "in_div_v_prev")):
            if ic.get(k_ic) is not None:
                parts[k_rec] = ic[k_ic]
    return parts


_PINNED_KEEP = []
_lib = None


def load_library():
    """dlopens the backend; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libswiftcuda.so is missing: run __graft_entry__.build() (no CPU fallback exists)")
        lib = C.CDLL(LIB_PATH)
        lib.swift_cuda_last_error.restype = C.c_char_p
        lib.swift_cuda_sizeof.restype = C.c_size_t
        lib.swift_cuda_halo_bytes_per_particle.restype = C.c_size_t
        lib.swift_cuda_halo_pack.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        lib.swift_cuda_halo_unpack.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        lib.swift_cuda_destroy.argtypes = [C.c_void_p]
        lib.swift_cuda_build_from_parts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        lib.swift_cuda_build_from_cells_parts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int,
                                                          C.c_void_p, C.c_int, C.c_void_p]
        lib.swift_cuda_step_parts.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.swift_cuda_set_max_active_bin.argtypes = [C.c_void_p, C.c_int]
        lib.swift_cuda_record_pairs.argtypes = [C.c_void_p, C.c_longlong]
        lib.swift_cuda_fetch_pairs.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p]
        lib.swift_cuda_set_time.argtypes = [C.c_void_p, C.c_longlong, C.c_int]
        lib.swift_cuda_drift.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p]
        lib.swift_cuda_kick.argtypes = [C.c_void_p, C.c_int]
        lib.swift_cuda_timestep.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.swift_cuda_reset_offsets.argtypes = [C.c_void_p, C.c_int]
        lib.swift_cuda_timestep_static.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        for name in ("swift_cuda_sphenix_set_props", "swift_cuda_sphenix_set_state", "swift_cuda_sphenix_get_results",
                     "swift_cuda_sphenix_get_counter"):
            getattr(lib, name).argtypes = [C.c_void_p, C.c_void_p]
        lib.swift_cuda_do_gradient.argtypes = [C.c_void_p]
        lib.swift_cuda_do_extra_ghost.argtypes = [C.c_void_p]
        lib.swift_cuda_eos_table_load_sesame.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.swift_cuda_eos_table_load_hm80.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p]
        lib.swift_cuda_eos_table_write_bin.argtypes = [C.c_char_p, C.c_void_p]
        lib.swift_cuda_eos_table_read_bin.argtypes = [C.c_char_p, C.c_void_p]
        lib.swift_cuda_eos_table_free.argtypes = [C.c_void_p]
        lib.swift_cuda_eos_table_free.restype = None
        lib.swift_cuda_particles_write_bin.argtypes = [C.c_char_p, C.c_void_p]
        lib.swift_cuda_particles_read_bin.argtypes = [C.c_char_p, C.c_void_p]
        lib.swift_cuda_particles_free.argtypes = [C.c_void_p]
        lib.swift_cuda_particles_free.restype = None
        lib.swift_cuda_set_eos_table.argtypes = [C.c_void_p, C.c_void_p]
        lib.swift_cuda_do_limiter_loop.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.swift_cuda_do_limiter.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.swift_cuda_get_limiter_timings.argtypes = [C.c_void_p, C.c_void_p]
        _lib = lib
    return _lib


class SwiftCudaError(RuntimeError):
    pass


def timestep_static(device, count, h, v_sig, h_dt, min_ngb_time_bin, time_bin, ti_current, max_active_bin, integ,
                    stream=None):
    """swift_cuda_timestep_static: runner_do_timestep on DEVICE arrays the caller owns (raw device pointers as
    ints), with the library's static scratch array: nothing is allocated.  integ: an Integrator.
    Returns (ti_end_min, updated); time_bin is updated in place."""
    lib = load_library()
    a = TimestepStaticArgs(count, h, v_sig, h_dt, min_ngb_time_bin, time_bin, ti_current, max_active_bin, integ)
    t, u = C.c_longlong(0), C.c_longlong(0)
    if lib.swift_cuda_timestep_static(device, C.byref(a), C.byref(t), C.byref(u), stream):
        raise SwiftCudaError(lib.swift_cuda_last_error().decode())
    return t.value, u.value


def host_plan(params, ic):
    """Host-only (no GPU): the tree, tasks and work-list size the library builds for `ic`."""
    lib = load_library()
    lib.swift_cuda_host_free.argtypes = [C.c_void_p]
    n = len(ic["h"])
    x = np.ascontiguousarray(ic["x"], dtype=np.float64)
    h = np.ascontiguousarray(ic["h"], dtype=np.float32)
    P = PartsIn()
    P.n = n
    P.x = x.ctypes.data_as(C.POINTER(C.c_double))
    P.h = h.ctypes.data_as(_FP)
    cells_p, tasks_p = C.POINTER(Cell)(), C.POINTER(Task)()
    nc, nt = C.c_int(), C.c_int()
    cdim = (C.c_int * 3)()
    ni, nitems = C.c_longlong(), C.c_longlong()
    order = np.zeros(n, dtype=np.int32)
    rc = lib.swift_cuda_host_plan(C.byref(params), C.byref(P), C.byref(cells_p), C.byref(nc), C.byref(tasks_p),
                                  C.byref(nt), order.ctypes.data_as(C.c_void_p), cdim, C.byref(ni), C.byref(nitems))
    if rc != 0:
        raise SwiftCudaError(lib.swift_cuda_last_error().decode())
    cells = (Cell * max(nc.value, 1))()
    C.memmove(cells, cells_p, C.sizeof(Cell) * nc.value)
    tasks = (Task * max(nt.value, 1))()
    C.memmove(tasks, tasks_p, C.sizeof(Task) * nt.value)
    lib.swift_cuda_host_free(cells_p)
    lib.swift_cuda_host_free(tasks_p)
    return {"cells": cells, "ncells": nc.value, "tasks": tasks, "ntasks": nt.value, "cdim": tuple(cdim),
            "order": order, "n_interactions": ni.value, "n_items": nitems.value}


def host_plan_digest(params, ic):
    """Host-only: a digest of everything the host builder makes for `ic` (swift_cuda_host_plan_digest)."""
    lib = load_library()
    keep = [np.ascontiguousarray(ic["x"], dtype=np.float64), np.ascontiguousarray(ic["h"], dtype=np.float32)]
    P = PartsIn()
    P.n = len(keep[1])
    P.x = keep[0].ctypes.data_as(C.POINTER(C.c_double))
    P.h = keep[1].ctypes.data_as(_FP)
    if ic.get("time_bin") is not None:
        keep.append(np.ascontiguousarray(ic["time_bin"], dtype=np.int8))
        P.time_bin = keep[-1].ctypes.data_as(C.POINTER(C.c_int8))
    d = C.c_ulonglong(0)
    if lib.swift_cuda_host_plan_digest(C.byref(params), C.byref(P), C.byref(d)) != 0:
        raise SwiftCudaError(lib.swift_cuda_last_error().decode())
    return d.value


def make_params(dim=(1.0, 1.0, 1.0), periodic=True, eta=1.2348, h_tolerance=1e-4, h_max=1e10, h_min=0.0,
                max_iter=30, alpha=0.8, max_active_bin=56, scheme=SCHEME_MINIMAL, til=(), idg_gamma=(5. / 3.,) * 4,
                fields_mode=0):
    p = Params()
    p.dim[:] = dim
    p.periodic = int(periodic)
    p.eta_neighbours = eta
    p.h_tolerance = h_tolerance
    p.h_max = h_max
    p.h_min = h_min
    p.max_smoothing_iterations = max_iter
    p.viscosity_alpha = alpha
    p.max_active_bin = max_active_bin
    p.scheme = scheme
    p.cosmo_a = 1.0
    p.cosmo_H = 0.0
    p.n_til = len(til)
    for k, t in enumerate(til):
        p.til[k] = t
    p.idg_gamma[:] = idg_gamma
    p.fields_mode = fields_mode
    return p


class HydroBackend:
    """One GPU's SPH neighbour-loop engine.  Usage:
        be = HydroBackend(params); be.build(ic); be.step(); out = be.results()
    """

    def __init__(self, params, device=0):
        self.lib = load_library()
        self.params = params
        self.ctx = C.c_void_p()
        self._keep = []
        self._check(self.lib.swift_cuda_create(C.byref(params), int(device), C.byref(self.ctx)))
        self.n = 0

    def _check(self, rc):
        if rc != 0:
            raise SwiftCudaError(self.lib.swift_cuda_last_error().decode())

    def close(self):
        if self.ctx:
            self.lib.swift_cuda_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _parts_in(self, ic):
        keep = []

        def arr(a, dt):
            a = np.ascontiguousarray(a, dtype=dt)
            keep.append(a)
            return a
        P = PartsIn()
        P.n = len(ic["h"])
        P.x = arr(ic["x"], np.float64).ctypes.data_as(C.POINTER(C.c_double))
        for k in ("v", "mass", "h", "u"):
            setattr(P, k, arr(ic[k], np.float32).ctypes.data_as(_FP))
        if ic.get("mat_id") is not None:
            P.mat_id = arr(ic["mat_id"], np.int32).ctypes.data_as(C.POINTER(C.c_int32))
        if ic.get("time_bin") is not None:
            P.time_bin = arr(ic["time_bin"], np.int8).ctypes.data_as(C.POINTER(C.c_int8))
        if ic.get("id") is not None:
            P.id = arr(ic["id"], np.int64).ctypes.data_as(C.POINTER(C.c_int64))
        for k in ("rho", "pressure", "f", "soundspeed", "balsara"):
            if ic.get("in_" + k) is not None:
                setattr(P, k, arr(ic["in_" + k], np.float32).ctypes.data_as(_FP))
        return P, keep

    # ---- building (role of space_rebuild + cache packing) ----
    def build(self, ic):
        P, keep = self._parts_in(ic)
        self.n = P.n
        self._check(self.lib.swift_cuda_build_from_particles(self.ctx, C.byref(P)))

    def build_from_cells(self, ic, cells, tasks, ntasks=None):
        """ic already in cell order; cells/tasks: ctypes arrays of Cell/Task (or layout-compatible)."""
        P, keep = self._parts_in(ic)
        self.n = P.n
        nt = len(tasks) if ntasks is None else ntasks
        self._check(self.lib.swift_cuda_build_from_cells(self.ctx, C.byref(P), len(cells), cells, nt, tasks))

    # ---- the caller's struct part records (the drop-in path of the C shim) ----
    def _layout(self):
        L = PartLayout()
        for k, v in PART_LAYOUTS[self.params.scheme].items():
            setattr(L, k, v)
        return L

    def build_from_parts(self, parts, cells=None, tasks=None, ntasks=None):
        """parts: array of part_dtype records.  cells/tasks given: the caller's tree (records in cell
        order, swift_cuda_build_from_cells_parts); otherwise the library bins them itself."""
        self.n = len(parts)
        L = self._layout()
        assert parts.dtype.itemsize == L.stride and parts.flags["C_CONTIGUOUS"]
        if cells is None:
            self._check(self.lib.swift_cuda_build_from_parts(self.ctx, C.byref(L), parts.ctypes.data_as(C.c_void_p),
                                                             C.c_int64(self.n)))
        else:
            nt = len(tasks) if ntasks is None else ntasks
            self._check(self.lib.swift_cuda_build_from_cells_parts(self.ctx, C.byref(L),
                                                                   parts.ctypes.data_as(C.c_void_p), C.c_int64(self.n),
                                                                   len(cells), cells, nt, tasks))

    def step_parts(self, parts, phases=PHASE_ALL):
        """Uploads the records, runs the phases, writes the results back into the records (in place)."""
        self._check(self.lib.swift_cuda_step_parts(self.ctx, parts.ctypes.data_as(C.c_void_p), int(phases)))

    def set_max_active_bin(self, max_active_bin):
        self._check(self.lib.swift_cuda_set_max_active_bin(self.ctx, int(max_active_bin)))

    def update_particles(self, ic):
        P, keep = self._parts_in(ic)
        self._check(self.lib.swift_cuda_update_particles(self.ctx, C.byref(P)))

    def set_transfer_count(self, count):
        """Only the first `count` particles (caller order) travel in update_particles / results / step_io."""
        self._check(self.lib.swift_cuda_set_transfer_count(self.ctx, C.c_int64(int(count))))

    def restore_particles(self):
        """Device-side reset of h to the values of the last build/update (no host traffic)."""
        self._check(self.lib.swift_cuda_restore_particles(self.ctx))

    # ---- operators ----
    def step(self, do_sort=True, restore=False):
        """sort -> density -> ghost -> force -> end_force.  restore: start from the h of the last build/update."""
        self._check(self.lib.swift_cuda_step(self.ctx, (1 if do_sort else 0) | (2 if restore else 0)))

    def step_io(self, ic, out, do_sort=True):
        """update_particles + step + results as one call, transfers overlapped with the kernels
        (swift_cuda_step_io); `ic` and `out` should live in page-locked host memory."""
        P, keep = self._parts_in(ic)
        O = PartsOut()
        for k, w in OUT_FIELDS + OUT_FIRST:
            if k in out:
                setattr(O, k, out[k].ctypes.data_as(_FP))
        O.min_ngb_time_bin = out["min_ngb_time_bin"].ctypes.data_as(C.POINTER(C.c_int8))
        self._check(self.lib.swift_cuda_step_io(self.ctx, C.byref(P), 1 if do_sort else 0, C.byref(O)))
        return out

    def measure_fp32_peak(self):
        self.lib.swift_cuda_measure_fp32_peak.restype = C.c_double
        self.lib.swift_cuda_measure_fp32_peak.argtypes = [C.c_void_p]
        return float(self.lib.swift_cuda_measure_fp32_peak(self.ctx))

    def runner_do_hydro_sort(self):
        self._check(self.lib.swift_cuda_do_sort(self.ctx))

    def runner_do_density(self):
        """All runner_do{self,pair}1_branch_density / runner_dosub_*1_density tasks."""
        self._check(self.lib.swift_cuda_do_density(self.ctx))

    def runner_do_ghost(self):
        self._check(self.lib.swift_cuda_do_ghost(self.ctx))

    def runner_do_force(self):
        """All runner_do{self,pair}2_branch_force tasks + runner_do_end_hydro_force."""
        self._check(self.lib.swift_cuda_do_force(self.ctx))

    # ---- SPHENIX: the gradient loop and the extra ghost ----
    def sphenix_set_props(self, viscosity_alpha_max=2.0, viscosity_alpha_min=0.0, viscosity_length=0.05,
                          diffusion_alpha=0.0, diffusion_beta=1.0, diffusion_alpha_max=1.0, diffusion_alpha_min=0.0,
                          time_base=1e-3, **_ignored):
        """hydro_props->viscosity / ->diffusion of the scheme and e->time_base; before the build."""
        p = SphenixProps(viscosity_alpha_max, viscosity_alpha_min, viscosity_length, diffusion_alpha, diffusion_beta,
                         diffusion_alpha_max, diffusion_alpha_min, time_base)
        self._check(self.lib.swift_cuda_sphenix_set_props(self.ctx, C.byref(p)))

    def sphenix_set_state(self, alpha_visc=None, alpha_diff=None, div_v_previous_step=None):
        """p->viscosity.alpha, p->diffusion.alpha, p->viscosity.div_v_previous_step (caller order); after the build."""
        S = SphenixIn()
        keep = []
        for k, a in (("alpha_visc", alpha_visc), ("alpha_diff", alpha_diff), ("div_v_previous_step", div_v_previous_step)):
            if a is not None:
                a = np.ascontiguousarray(a, dtype=np.float32)
                assert len(a) == self.n
                keep.append(a)
                setattr(S, k, a.ctypes.data_as(_FP))
        self._check(self.lib.swift_cuda_sphenix_set_state(self.ctx, C.byref(S)))

    def sphenix_results(self):
        out = {k: np.zeros(self.n, dtype=np.float32) for k in SPHENIX_OUT_FIELDS}
        O = SphenixOut()
        for k in SPHENIX_OUT_FIELDS:
            setattr(O, k, out[k].ctypes.data_as(_FP))
        self._check(self.lib.swift_cuda_sphenix_get_results(self.ctx, C.byref(O)))
        n = C.c_longlong(0)
        self._check(self.lib.swift_cuda_sphenix_get_counter(self.ctx, C.byref(n)))
        out["n_gradient"] = n.value
        return out

    def runner_do_gradient(self):
        """All runner_do{self,pair}1_branch_gradient / runner_dosub_*1_gradient tasks."""
        self._check(self.lib.swift_cuda_do_gradient(self.ctx))

    def runner_do_extra_ghost(self):
        self._check(self.lib.swift_cuda_do_extra_ghost(self.ctx))

    # ---- time integration: the particles stay on the device across steps ----
    def integrator_init(self, cfl_condition, log_max_h_change, minimal_internal_energy, dt_min, dt_max, time_base,
                        **_ignored):
        """hydro_first_init_part on the device copy; the constants of e->hydro_properties / e->dt_* / e->time_base."""
        p = Integrator(cfl_condition, log_max_h_change, minimal_internal_energy, dt_min, dt_max, time_base)
        self._check(self.lib.swift_cuda_integrator_init(self.ctx, C.byref(p)))

    def set_time(self, ti_current, max_active_bin):
        self._check(self.lib.swift_cuda_set_time(self.ctx, ti_current, max_active_bin))

    def drift(self, ti_old):
        """runner_do_drift_part on every particle up to the current time; returns (dx_max_part, dx_max_sort)."""
        dx = (C.c_float * 2)()
        self._check(self.lib.swift_cuda_drift(self.ctx, ti_old, dx))
        return float(dx[0]), float(dx[1])

    def kick(self, which):
        """which = 1: runner_do_kick1, 2: runner_do_kick2 (+ hydro_reset_predicted_values)."""
        self._check(self.lib.swift_cuda_kick(self.ctx, which))

    def timestep(self):
        """runner_do_timestep; returns (ti_end_min, updated)."""
        t, u = C.c_longlong(0), C.c_longlong(0)
        self._check(self.lib.swift_cuda_timestep(self.ctx, C.byref(t), C.byref(u)))
        return t.value, u.value

    # ---- table equations of state (Planetary scheme) ----
    def set_eos_table(self, table):
        """Copies an EosTable (load_eos_table) to the device: particles of its mat_id are evaluated from it."""
        self._check(self.lib.swift_cuda_set_eos_table(self.ctx, C.byref(table)))

    # ---- time-step limiter (swift --limiter) ----
    def runner_do_limiter_loop(self, time_bin=None, want_wakeup=False):
        """The tasks of subtype limiter (runner_do{self,pair}1_branch_limiter, runner_dosub_*1_limiter) over the
        whole box, after the time-step task.  time_bin: the new bins of all particles (caller order) when the
        caller runs runner_do_timestep itself; None in the time-integration mode.  Returns
        p->limiter_data.wakeup (int8, -56 = not awake) when want_wakeup."""
        tb = None
        if time_bin is not None:
            tb = np.ascontiguousarray(time_bin, dtype=np.int8)
            assert len(tb) == self.n
        wake = np.zeros(self.n, dtype=np.int8) if want_wakeup else None
        self._check(self.lib.swift_cuda_do_limiter_loop(self.ctx, tb.ctypes.data if tb is not None else None,
                                                        wake.ctypes.data if wake is not None else None))
        return wake

    def runner_do_limiter(self):
        """runner_do_limiter on the device state; returns (earliest new end of step, woken particles)."""
        t, u = C.c_longlong(0), C.c_longlong(0)
        self._check(self.lib.swift_cuda_do_limiter(self.ctx, C.byref(t), C.byref(u)))
        return t.value, u.value

    def limiter_timings(self):
        ms = (C.c_float * 2)()
        self._check(self.lib.swift_cuda_get_limiter_timings(self.ctx, ms))
        return {"limiter_loop_ms": ms[0], "limiter_ms": ms[1]}

    def integrator_timings(self):
        ms = (C.c_float * 4)()
        self._check(self.lib.swift_cuda_get_integrator_timings(self.ctx, ms))
        return {"drift_ms": ms[0], "kick1_ms": ms[1], "kick2_ms": ms[2], "timestep_ms": ms[3]}

    def reset_offsets(self, what=1):
        self._check(self.lib.swift_cuda_reset_offsets(self.ctx, what))

    def state(self):
        """Every per-particle field of the device state, caller order."""
        n = self.n
        out = {k: np.zeros((n, w) if w > 1 else n, dtype=np.float64 if k == "x" else np.float32)
               for k, w in STATE_FIELDS}
        out["time_bin"] = np.zeros(n, dtype=np.int8)
        out["min_ngb_time_bin"] = np.zeros(n, dtype=np.int8)
        S = StateOut()
        S.x = out["x"].ctypes.data_as(C.POINTER(C.c_double))
        for k, _ in STATE_FIELDS[1:]:
            setattr(S, k, out[k].ctypes.data_as(_FP))
        S.time_bin = out["time_bin"].ctypes.data_as(C.POINTER(C.c_int8))
        S.min_ngb_time_bin = out["min_ngb_time_bin"].ctypes.data_as(C.POINTER(C.c_int8))
        self._check(self.lib.swift_cuda_get_state(self.ctx, C.byref(S)))
        return out

    # ---- results ----
    def alloc_results(self, first_pass=False, pinned=False):
        """Output arrays for results(out=...); pinned=True allocates page-locked host memory (torch)."""
        n = self.n
        fields = OUT_FIELDS + (OUT_FIRST if first_pass else []) + [("min_ngb_time_bin", 1)]
        out = {}
        for k, w in fields:
            dt = np.int8 if k == "min_ngb_time_bin" else np.float32
            shape = (n, w) if w > 1 else (n,)
            if pinned:
                import torch
                t = torch.empty(shape, dtype=torch.int8 if dt == np.int8 else torch.float32, pin_memory=True)
                out[k] = t.numpy()
                self._keep.append(t)
            else:
                out[k] = np.zeros(shape, dtype=dt)
        return out

    def results(self, first_pass=False, out=None):
        if out is None:
            out = self.alloc_results(first_pass)
        P = PartsOut()
        for k, w in OUT_FIELDS + OUT_FIRST:
            if k in out:
                setattr(P, k, out[k].ctypes.data_as(_FP))
        P.min_ngb_time_bin = out["min_ngb_time_bin"].ctypes.data_as(C.POINTER(C.c_int8))
        self._check(self.lib.swift_cuda_get_results(self.ctx, C.byref(P)))
        return out

    def counters(self):
        c = Counters()
        self._check(self.lib.swift_cuda_get_counters(self.ctx, C.byref(c)))
        return {k: getattr(c, k) for k, _ in Counters._fields_}

    def timings(self):
        t = Timings()
        self._check(self.lib.swift_cuda_get_timings(self.ctx, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timings._fields_}

    def tree(self):
        nc, nt = C.c_int(), C.c_int()
        cdim = (C.c_int * 3)()
        self._check(self.lib.swift_cuda_get_tree_size(self.ctx, C.byref(nc), C.byref(nt), cdim))
        cells = (Cell * max(nc.value, 1))()
        tasks = (Task * max(nt.value, 1))()
        order = np.zeros(self.n, dtype=np.int32)
        self._check(self.lib.swift_cuda_get_tree(self.ctx, cells, tasks, order.ctypes.data_as(C.c_void_p)))
        return {"cells": cells, "ncells": nc.value, "tasks": tasks, "ntasks": nt.value, "cdim": tuple(cdim),
                "order": order}

    # ---- multi-GPU support ----
    def touched_mask(self):
        m = np.zeros(self.n, dtype=np.uint8)
        self._check(self.lib.swift_cuda_get_touched(self.ctx, m.ctypes.data_as(C.c_void_p)))
        return m.astype(bool)

    def refresh_cell_hmax(self, after_ghost):
        self._check(self.lib.swift_cuda_refresh_cell_hmax(self.ctx, int(after_ghost)))

    def halo_bytes(self, phase):
        return int(self.lib.swift_cuda_halo_bytes_per_particle(int(phase)))

    def halo_pack(self, phase, index_ptr, count, buffer_ptr, stream=0):
        self._check(self.lib.swift_cuda_halo_pack(self.ctx, int(phase), C.c_void_p(index_ptr), C.c_int64(count),
                                                  C.c_void_p(buffer_ptr), C.c_void_p(stream)))

    def halo_unpack(self, phase, index_ptr, count, buffer_ptr, stream=0):
        self._check(self.lib.swift_cuda_halo_unpack(self.ctx, int(phase), C.c_void_p(index_ptr), C.c_int64(count),
                                                    C.c_void_p(buffer_ptr), C.c_void_p(stream)))

    # ---- test hook ----
    def record_pairs(self, capacity):
        self._check(self.lib.swift_cuda_record_pairs(self.ctx, int(capacity)))

    def fetch_pairs(self, loop, capacity):
        cnt = C.c_longlong()
        self._check(self.lib.swift_cuda_fetch_pairs(self.ctx, loop, None, 0, C.byref(cnt)))
        buf = np.zeros((max(cnt.value, 1), 2), dtype=np.int64)
        self._check(self.lib.swift_cuda_fetch_pairs(self.ctx, loop, buf.ctypes.data_as(C.c_void_p), cnt.value,
                                                    C.byref(cnt)))
        return buf[:cnt.value]
