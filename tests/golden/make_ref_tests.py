"""Runs the UNMODIFIED reference unit tests (tests/test27cells.c, tests/test125cells.c, tests/testPeriodicBC.c, linked against
the reference's own runner_doiact_hydro.o by oracle/Makefile.shim, target `cpu`) in this container and
stores their output tables as fixtures: what tests/test_reference_tests.py compares the same tests,
linked against the shim + libswiftcuda.so instead, with on the GPU box."""
import os
import shutil
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "..", "oracle", "_ref")
RUNS = {
    "test27cells": (["-n", "4", "-r", "1", "-d", "0", "-f", "standard"], "swift_dopair_27_standard.dat"),
    "test27cells_perturbed": (["-n", "4", "-r", "1", "-d", "0.1", "-f", "perturbed"], "swift_dopair_27_perturbed.dat"),
    "test125cells": (["-n", "6", "-r", "1", "-v", "0", "-p", "0", "-f", "standard"], "swift_dopair_125_standard.dat"),
    "test125cells_v2": (["-n", "6", "-r", "1", "-v", "2", "-p", "1", "-f", "standard"], "swift_dopair_125_standard.dat"),
}
# the same four runs with the reference compiled as SPHENIX_SPH (make -f Makefile.shim VARIANT=sphenix cpu): the
# gradient loop and the extra ghost are part of test125cells then (tests/test125cells.c:724-767)
for variant, suffix, prefix in (("minimal", "", ""), ("sphenix", "_sphenix", "sphenix_")):
    for name, (args, out) in RUNS.items():
        exe = os.path.join(REF, name.split("_")[0] + "_ref" + suffix)
        if not os.path.exists(exe):
            print("skip", variant, name, "(not built)")
            continue
        with tempfile.TemporaryDirectory() as tmp:
            subprocess.run([exe] + args, cwd=tmp, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            shutil.copy(os.path.join(tmp, out), os.path.join(HERE, "ref_tests", prefix + name + ".dat"))
            print(variant, name, "->", out)

# tests/testPeriodicBC.c (26 boundary cells of an 8^3 periodic grid, each against its 26 wrapped neighbours; the
# reference's script runs -v 0..3, tests/testPeriodicBC.sh.in:3-11): the 1 MB tables are kept as compressed arrays
import numpy as np  # noqa: E402
PBC = {"testPeriodicBC_v0": ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "0"],
       "testPeriodicBC_v2": ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "2"]}
for variant, suffix, prefix in (("minimal", "", ""), ("sphenix", "_sphenix", "sphenix_")):
    for name, args in PBC.items():
        exe = os.path.join(REF, "testPeriodicBC_ref" + suffix)
        if not os.path.exists(exe):
            print("skip", variant, name, "(not built)")
            continue
        with tempfile.TemporaryDirectory() as tmp:
            subprocess.run([exe] + args, cwd=tmp, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            rows = [l.split() for l in open(os.path.join(tmp, "swift_periodic_BC_standard.dat"))
                    if l.strip() and not l.lstrip().startswith("#")]
            np.savez_compressed(os.path.join(HERE, "ref_tests", prefix + name + ".npz"), table=np.array(rows, dtype=np.float64))
            print(variant, name, "->", len(rows), "rows")

# tests/testActivePair.c: two neighbouring cells with every mix of active and inactive particles (one particle per
# cell, all active, half, none, 10 %, one cell active and one not, ...), the density pair task and the force pair
# task against the brute force of tests/tools.c; the reference's script (tests/testActivePair.sh.in) runs a random
# seed and the seed 1506434777 "that triggered a bug"; fixed seeds here
ACTIVE_PAIR = {"testActivePair_s1506434777": "1506434777", "testActivePair_s12345": "12345"}
for variant, suffix, prefix in (("minimal", "", ""), ("sphenix", "_sphenix", "sphenix_")):
    exe = os.path.join(REF, "testActivePair_ref" + suffix)
    if not os.path.exists(exe):
        print("skip", variant, "testActivePair (not built)")
        continue
    for name, seed in ACTIVE_PAIR.items():
        with tempfile.TemporaryDirectory() as tmp:
            subprocess.run([exe, "-n", "6", "-r", "1", "-d", "0", "-f", "active", "-s", seed], cwd=tmp, check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            z = {}
            for key, fname in (("density", "swift_dopair_active.dat"), ("force", "swift_dopair2_force_active.dat")):
                rows = [l.split() for l in open(os.path.join(tmp, fname)) if l.strip() and not l.lstrip().startswith("#")]
                z[key] = np.array(rows, dtype=np.float64)
            np.savez_compressed(os.path.join(HERE, "ref_tests", prefix + name + ".npz"), **z)
            print(variant, name, "->", {k: v.shape for k, v in z.items()})

# The reference's PERTURBED runs with its scripts' own arguments (tests/test27cellsPerturbed.sh.in,
# test125cellsPerturbed.sh.in, testPeriodicBCPerturbed.sh.in): positions jittered by 0.1 spacings, smoothing lengths
# spread by up to 10 % / 30 % (-p 1.1 / 1.3).  Tables as compressed arrays (first table each run writes).
PERTURBED = {
    "test27cells_n6_perturbed": ("test27cells", ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "2"],
                                 "swift_dopair_27_perturbed.dat"),
    "test27cells_n6_perturbed_h": ("test27cells", ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "1", "-p", "1.1"],
                                   "swift_dopair_27_perturbed.dat"),
    "test27cells_n6_perturbed_h2": ("test27cells", ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "3", "-p", "1.3"],
                                    "swift_dopair_27_perturbed.dat"),
    "test125cells_perturbed": ("test125cells", ["-n", "6", "-r", "1", "-d", "0.1", "-v", "1", "-p", "2", "-f", "perturbed"],
                               "swift_dopair_125_perturbed.dat"),
    "testPeriodicBC_perturbed": ("testPeriodicBC", ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "1"],
                                 "swift_periodic_BC_perturbed.dat"),
    # test27cells_subset (tests/Makefile.am:97-99): the same test with runner_do{self,pair}_subset_branch_density in
    # the place of the density tasks (tests/test27cells.sh.in runs both binaries with the same arguments)
    "test27cells_subset_n6": ("test27cells_subset", ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "2"],
                              "swift_dopair_27_standard.dat"),
    "test27cells_subset_n6_h": ("test27cells_subset", ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "1", "-p", "1.1"],
                                "swift_dopair_27_standard.dat"),
}
for variant, suffix, prefix in (("minimal", "", ""), ("sphenix", "_sphenix", "sphenix_")):
    for name, (exe_name, args, out) in PERTURBED.items():
        exe = os.path.join(REF, exe_name + "_ref" + suffix)
        if not os.path.exists(exe):
            print("skip", variant, name, "(not built)")
            continue
        with tempfile.TemporaryDirectory() as tmp:
            subprocess.run([exe] + args, cwd=tmp, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            rows = [l.split() for l in open(os.path.join(tmp, out)) if l.strip() and not l.lstrip().startswith("#")]
            np.savez_compressed(os.path.join(HERE, "ref_tests", prefix + name + ".npz"), table=np.array(rows, dtype=np.float64))
            print(variant, name, "->", len(rows), "rows")
