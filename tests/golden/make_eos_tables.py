"""Writes tests/golden/eos_tables/: SYNTHETIC equation-of-state tables in the reference's two text formats
(src/equation_of_state/planetary/sesame.h:133-148 and hm80.h:78-91), and what the REFERENCE's own loaders and
interpolation functions (oracle/_ref/libswiftref_planetary.so) make of them: the prepared arrays and P, c on a
grid of (rho, u) points.

The reference's real table files are not part of its tree (examples/Planetary/EoSTables/get_eos_tables.sh downloads
them; no network here), so these tables are NOT physical data: they are smooth (a cold curve plus a thermal term)
and salted with what the reference's code has branches for -- a duplicate density, u slices that are not
monotonic at low T, non-positive u, P and c entries (tension at low density).  Run in the build container:

    python tests/golden/make_eos_tables.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
OUT_DIR = os.path.join(HERE, "eos_tables")

# name: (file, format, mat_id)
TABLES = {
    "sesame_like_iron": ("sesame_like_iron.txt", "sesame", 300),
    "aneos_like_forsterite": ("aneos_like_forsterite.txt", "sesame", 400),
    "custom_0": ("custom_0.txt", "sesame", 900),
    "hm80_like_HHe": ("hm80_like_HHe.txt", "hm80", 200),
    "hm80_like_rock": ("hm80_like_rock.txt", "hm80", 202),
}


def write_sesame(path, rho0, cv, seed, num_rho=33, num_T=25, date=20220714):
    """SI units.  The first density and temperature (0) are what the reference drops."""
    rng = np.random.default_rng(seed)
    rho = np.concatenate([[0.0], np.geomspace(rho0 * 2e-3, rho0 * 40.0, num_rho - 1)])
    rho[10] = rho[9]                                     # a duplicate density (sesame.h:520-525)
    T = np.concatenate([[0.0], np.geomspace(20.0, 2e6, num_T - 1)])
    lines = ["# synthetic SESAME-style table for the parity tests of swift_subtask_dev_b200 (not physical data)"]
    lines += ["# header line %d" % k for k in range(2, 13)]
    lines.append("%d" % date)
    lines.append("%d %d" % (num_rho, num_T))
    lines.append(" ".join("%.8e" % v for v in rho))
    lines.append(" ".join("%.8e" % v for v in T))
    for i_T in range(num_T):
        for i_rho in range(num_rho):
            r, t = rho[i_rho], T[i_T]
            x = r / rho0 if r > 0 else 0.0
            u_cold = 2e5 * (x ** 0.7 - 1.0) if r > 0 else 0.0        # negative at low density
            u = u_cold + cv * t * (1.0 + 0.2 * np.sin(3.0 * np.log(x + 1e-9)))
            if 0 < i_T < 5 and i_rho > num_rho - 8:
                u *= 1.0 + 0.05 * (4 - i_T) * rng.random()          # slices that fall with T at high density
            P = 1.2e11 * (x ** 2.3 - x ** 1.1) * x ** 2 + 0.6 * r * cv * t   # tension for 0 < x < 1 at low T
            c2 = 1.2e11 / rho0 * (2.3 * x ** 1.3 - 1.1 * x ** 0.1) + 1.0 * cv * t if r > 0 else 0.0
            c = np.sqrt(c2) if c2 > 0 else 0.0
            s = cv * np.log1p(t / 50.0) * (1.0 + 0.1 * x) if t > 0 else 0.0
            lines.append("%.8e %.8e %.8e %.8e" % (u, P, c, s))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def write_hm80(path, seed, num_rho=41, num_u=31, date=20201003):
    rng = np.random.default_rng(seed)
    log_rho_min, log_rho_max = np.log(1e-1), np.log(2e4)
    log_u_min, log_u_max = np.log(1e4), np.log(1e9)
    lines = ["# synthetic HM80-style table for the parity tests of swift_subtask_dev_b200 (not physical data)"]
    lines += ["# header line %d" % k for k in range(2, 6)]
    lines.append("%d" % date)
    lines.append("%.8e %.8e %d %.8e %.8e %d" % (log_rho_min, log_rho_max, num_rho, log_u_min, log_u_max, num_u))
    lr = np.linspace(log_rho_min, log_rho_max, num_rho)
    lu = np.linspace(log_u_min, log_u_max, num_u)
    for i in range(num_rho):
        rho = np.exp(lr[i])
        row = 0.55 * rho * np.exp(lu) * (1.0 + 0.3 * np.tanh(lr[i] - 5.0)) + 3e4 * rho ** 1.8
        row *= 1.0 + 0.02 * rng.random(num_u)
        lines.append(" ".join("%.8e" % v for v in row))
    for i in range(num_rho):                                          # temperatures: never read by the reference
        lines.append(" ".join("%.8e" % (np.exp(lu[j]) / 1.2e4) for j in range(num_u)))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def main():
    import eos_lib as el
    os.makedirs(OUT_DIR, exist_ok=True)
    write_sesame(os.path.join(OUT_DIR, "sesame_like_iron.txt"), 7.9e3, 450.0, 1)
    write_sesame(os.path.join(OUT_DIR, "aneos_like_forsterite.txt"), 3.3e3, 1100.0, 2, num_rho=29, num_T=21)
    write_sesame(os.path.join(OUT_DIR, "custom_0.txt"), 1.0e3, 4000.0, 3, num_rho=17, num_T=13, date=19990101)
    write_hm80(os.path.join(OUT_DIR, "hm80_like_HHe.txt"), 4)
    write_hm80(os.path.join(OUT_DIR, "hm80_like_rock.txt"), 5, num_rho=25, num_u=19)
    units = el.ref_units()
    z = {"units": np.array([units.density, units.energy_per_unit_mass, units.pressure, units.speed])}
    for name, (fname, fmt, mat_id) in TABLES.items():
        t = el.ref_load(mat_id, os.path.join(OUT_DIR, fname))
        a = el.table_arrays(t)
        rho, u = el.sample_points(a, units.density)
        P, c = el.ref_eval(mat_id, rho, u)
        for k, v in a.items():
            z["%s/table/%s" % (name, k)] = np.asarray(v)
        z[name + "/rho"], z[name + "/u"], z[name + "/P"], z[name + "/c"] = rho, u, P, c
        print(name, "mat_id", mat_id, "table", a["num_rho"], "x", a["num_T"], "points", len(rho),
              "P zero", int((P == 0).sum()), "P range %.3g..%.3g" % (P[P > 0].min(), P.max()),
              "c range %.3g..%.3g" % (c.min(), c.max()), "nan", int(np.isnan(P).sum() + np.isnan(c).sum()))
    np.savez_compressed(os.path.join(OUT_DIR, "reference_results.npz"), **z)


if __name__ == "__main__":
    main()
