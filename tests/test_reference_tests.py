"""The reference's OWN unit tests (tests/test27cells.c, tests/test125cells.c, tests/testPeriodicBC.c, tests/testActivePair.c), compiled from where
they lie and linked against swift_subtask_dev_b200/shim/swift_cuda_shim.c + libswiftcuda.so instead
of the reference's runner_doiact_hydro.o (oracle/Makefile.shim).  Their output tables are gated the
way the reference gates them (tests/difffloat.py:75-130 with tests/tolerance_*.dat: absolute,
relative and "too small to compare" thresholds per column) against the brute-force table the test
itself computes, and compared with the table the unmodified reference wrote for the same arguments
(tests/golden/ref_tests, tests/golden/make_ref_tests.py)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
GOLD = os.path.join(ROOT, "tests", "golden", "ref_tests")


def load_table(path):
    rows = [l.split() for l in open(path) if l.strip() and not l.lstrip().startswith("#")]
    return np.array(rows, dtype=np.float64)


def difffloat(a, b, tol, number_to_check):
    """The semantics of the reference's tests/difffloat.py:83-127 (the reference's scripts pass
    number_to_check = 6: only the first n^3 rows, the main cell, are gated); returns the failing
    (row, column)s."""
    abs_tol, rel_tol, lim_tol = tol[0], tol[1], tol[2]
    assert a.shape == b.shape
    nrows = min(number_to_check ** 3, len(a))
    a, b = a[:nrows], b[:nrows]
    diff = np.abs(a - b)
    s = np.abs(a + b)
    rel = np.where(s > 0, diff / np.where(s > 0, s, 1.0), 0.0)
    bad = diff > 1.1 * abs_tol[None, :]
    small = (np.abs(a) + np.abs(b)) < lim_tol[None, :]
    bad |= (~small) & (rel > 1.1 * rel_tol[None, :])
    return np.argwhere(bad)


def run_test(exe, args, tmp_path, timeout=900):
    path = os.path.join(REF, exe)
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/%s not built (make -C oracle -f Makefile.shim needs the reference tree)" % exe)
    r = subprocess.run([path] + args, cwd=tmp_path, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r


CASES_27 = [("test27cells", ["-n", "4", "-r", "1", "-d", "0", "-f", "standard"], "standard", "tolerance_27_normal.dat"),
            ("test27cells_perturbed", ["-n", "4", "-r", "1", "-d", "0.1", "-f", "perturbed"], "perturbed",
             "tolerance_27_perturbed.dat")]


# The same tests with the reference compiled as SPHENIX_SPH, its configure default (Makefile.shim VARIANT=sphenix):
# test125cells then runs runner_do{self,pair}1_branch_gradient and runner_do_extra_ghost between the ghost and the
# force loop (tests/test125cells.c:724-767), all of them the shim's; the record is the scheme's 160-byte struct part.
VARIANTS = [("", ""), ("_sphenix", "sphenix_")]


@pytest.mark.gpu
@pytest.mark.parametrize("suffix,prefix", VARIANTS, ids=["minimal", "sphenix"])
@pytest.mark.parametrize("name,args,ext,tolfile", CASES_27)
def test_test27cells_through_the_shim(name, args, ext, tolfile, suffix, prefix, tmp_path):
    run_test("test27cells_cuda" + suffix, args, tmp_path)
    ours = load_table(tmp_path / ("swift_dopair_27_%s.dat" % ext))
    brute = load_table(tmp_path / ("brute_force_27_%s.dat" % ext))
    tol = load_table(os.path.join(REF, tolfile))
    bad = difffloat(ours, brute, tol, int(args[1]))
    assert len(bad) == 0, "difffloat gate: %d entries out of tolerance, first (row, col) %s" % (len(bad), bad[:5])
    # and against the unmodified reference's own table: the per-pair partial sums are added in the
    # same order; only the grouping of the additions across calls differs
    ref = load_table(os.path.join(GOLD, prefix + name + ".dat"))
    bad = difffloat(ours, ref, tol, 10 ** 6)  # the reference's gate on ALL rows (the tables print 7 digits)
    assert len(bad) == 0, "against the reference's own table: %d entries, first (row, col) %s" % (len(bad), bad[:5])


CASES_125 = [("test125cells", ["-n", "6", "-r", "1", "-v", "0", "-p", "0", "-f", "standard"], "tolerance_125_normal.dat"),
             ("test125cells_v2", ["-n", "6", "-r", "1", "-v", "2", "-p", "1", "-f", "standard"], "tolerance_125_normal.dat")]


@pytest.mark.gpu
@pytest.mark.parametrize("suffix,prefix", VARIANTS, ids=["minimal", "sphenix"])
@pytest.mark.parametrize("name,args,tolfile", CASES_125)
def test_test125cells_through_the_shim(name, args, tolfile, suffix, prefix, tmp_path):
    run_test("test125cells_cuda" + suffix, args, tmp_path)
    ours = load_table(tmp_path / "swift_dopair_125_standard.dat")
    brute = load_table(tmp_path / "brute_force_125_standard.dat")
    tol = load_table(os.path.join(REF, tolfile))
    bad = difffloat(ours, brute, tol, int(args[1]))
    assert len(bad) == 0, "difffloat gate: %d entries out of tolerance, first (row, col) %s" % (len(bad), bad[:5])
    ref = load_table(os.path.join(GOLD, prefix + name + ".dat"))
    bad = difffloat(ours, ref, tol, 10 ** 6)  # the reference's gate on ALL rows (the tables print 7 digits)
    assert len(bad) == 0, "against the reference's own table: %d entries, first (row, col) %s" % (len(bad), bad[:5])


# tests/testPeriodicBC.c: an 8^3 grid of cells in a periodic box; for each of 26 boundary cells (corners, face
# and edge centres) the density pair task against its 26 neighbours THROUGH the box wrap and the self task
# (tests/testPeriodicBC.c:295-336, 545-627).  The reference's script (tests/testPeriodicBC.sh.in:3-11) gates the
# first 6^3 rows with tolerance_periodic_BC_normal.dat for -v 0..3; here every row of the table is gated.
CASES_PBC = [("testPeriodicBC_v0", ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "0"]),
             ("testPeriodicBC_v2", ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "2"])]


@pytest.mark.gpu
@pytest.mark.parametrize("suffix,prefix", VARIANTS, ids=["minimal", "sphenix"])
@pytest.mark.parametrize("name,args", CASES_PBC)
def test_testPeriodicBC_through_the_shim(name, args, suffix, prefix, tmp_path):
    run_test("testPeriodicBC_cuda" + suffix, args, tmp_path)
    ours = load_table(tmp_path / "swift_periodic_BC_standard.dat")
    brute = load_table(tmp_path / "brute_force_periodic_BC_standard.dat")
    tol = load_table(os.path.join(REF, "tolerance_periodic_BC_normal.dat"))
    assert len(ours) == 26 * 6 ** 3
    bad = difffloat(ours, brute, tol, 10 ** 6)
    assert len(bad) == 0, "difffloat gate: %d entries out of tolerance, first (row, col) %s" % (len(bad), bad[:5])
    ref = np.load(os.path.join(GOLD, prefix + name + ".npz"))["table"]
    bad = difffloat(ours, ref, tol, 10 ** 6)
    assert len(bad) == 0, "against the reference's own table: %d entries, first (row, col) %s" % (len(bad), bad[:5])


# tests/testActivePair.c: two neighbouring cells in every mix of active and inactive particles (one particle per cell;
# all active; half; none; 10 %; one cell active, the other not; ...), first the density pair task, then the force pair
# task, each against the brute force of tests/tools.c (tests/testActivePair.c:343-560).  The reference's script
# (tests/testActivePair.sh.in) gates every row with tolerance_pair_active.dat and tolerance_pair_force_active.dat, for
# a random seed and for the seed 1506434777 "that triggered a bug"; fixed seeds here.
# NOT YET RUN ON HARDWARE (linked after the round's GPU budget was spent; the per-task entries it calls are the ones
# test27cells / test125cells / testPeriodicBC pass through, the mixes of inactive particles are new): xfail until seen.
CASES_ACTIVE_PAIR = [("testActivePair_s1506434777", "1506434777"), ("testActivePair_s12345", "12345")]


@pytest.mark.gpu
@pytest.mark.xfail(strict=False, reason="first run on hardware (linked after the round's GPU budget was spent)")
@pytest.mark.parametrize("suffix,prefix", VARIANTS, ids=["minimal", "sphenix"])
@pytest.mark.parametrize("name,seed", CASES_ACTIVE_PAIR)
def test_testActivePair_through_the_shim(name, seed, suffix, prefix, tmp_path):
    run_test("testActivePair_cuda" + suffix, ["-n", "6", "-r", "1", "-d", "0", "-f", "active", "-s", seed], tmp_path,
             timeout=90)
    ref = np.load(os.path.join(GOLD, prefix + name + ".npz"))
    for key, ours_f, brute_f, tol_f in (
            ("density", "swift_dopair_active.dat", "brute_force_pair_active.dat", "tolerance_pair_active.dat"),
            ("force", "swift_dopair2_force_active.dat", "brute_force_dopair2_active.dat",
             "tolerance_pair_force_active.dat")):
        ours = load_table(tmp_path / ours_f)
        brute = load_table(tmp_path / brute_f)
        tol = load_table(os.path.join(REF, tol_f))
        bad = difffloat(ours, brute, tol, 10 ** 6)
        assert len(bad) == 0, "%s, difffloat gate: %d entries out of tolerance, first (row, col) %s" % (key, len(bad), bad[:5])
        bad = difffloat(ours, ref[key], tol, 10 ** 6)
        assert len(bad) == 0, "%s, against the reference's own table: %d entries, first %s" % (key, len(bad), bad[:5])


# The reference's PERTURBED runs with its scripts' own arguments (tests/test27cellsPerturbed.sh.in,
# test125cellsPerturbed.sh.in, testPeriodicBCPerturbed.sh.in): positions jittered by 0.1 spacings; smoothing lengths
# spread by up to 10 % and 30 % (-p 1.1 / 1.3, gated with tolerance_27_perturbed_h / _h2.dat).
# NOT YET RUN ON HARDWARE with these arguments (added after the round's GPU budget was spent; the binaries and the
# entries are those of the tests above): xfail until seen.
CASES_PERTURBED = [
    ("test27cells_n6_perturbed", "test27cells", ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "2"],
     "swift_dopair_27_perturbed.dat", "brute_force_27_perturbed.dat", "tolerance_27_perturbed.dat", 6),
    ("test27cells_n6_perturbed_h", "test27cells",
     ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "1", "-p", "1.1"],
     "swift_dopair_27_perturbed.dat", "brute_force_27_perturbed.dat", "tolerance_27_perturbed_h.dat", 6),
    ("test27cells_n6_perturbed_h2", "test27cells",
     ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "3", "-p", "1.3"],
     "swift_dopair_27_perturbed.dat", "brute_force_27_perturbed.dat", "tolerance_27_perturbed_h2.dat", 6),
    ("test125cells_perturbed", "test125cells", ["-n", "6", "-r", "1", "-d", "0.1", "-v", "1", "-p", "2", "-f", "perturbed"],
     "swift_dopair_125_perturbed.dat", "brute_force_125_perturbed.dat", "tolerance_125_perturbed.dat", 6),
    ("testPeriodicBC_perturbed", "testPeriodicBC", ["-n", "6", "-r", "1", "-d", "0.1", "-f", "perturbed", "-v", "1"],
     "swift_periodic_BC_perturbed.dat", "brute_force_periodic_BC_perturbed.dat", "tolerance_periodic_BC_perturbed.dat",
     10 ** 6),
    # test27cells_subset (tests/Makefile.am:97-99, tests/test27cells.sh.in): test27cells.c with
    # runner_do{self,pair}_subset_branch_density in the place of the density tasks; the shim runs them as the density
    # task with nobody but the subset active
    ("test27cells_subset_n6", "test27cells_subset", ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "2"],
     "swift_dopair_27_standard.dat", "brute_force_27_standard.dat", "tolerance_27_normal.dat", 6),
    ("test27cells_subset_n6_h", "test27cells_subset",
     ["-n", "6", "-r", "1", "-d", "0", "-f", "standard", "-v", "1", "-p", "1.1"],
     "swift_dopair_27_standard.dat", "brute_force_27_standard.dat", "tolerance_27_perturbed_h.dat", 6),
]


@pytest.mark.gpu
@pytest.mark.xfail(strict=False, reason="first run on hardware with these arguments (added after the round's GPU budget was spent)")
@pytest.mark.parametrize("suffix,prefix", VARIANTS, ids=["minimal", "sphenix"])
@pytest.mark.parametrize("name,exe,args,ours_f,brute_f,tol_f,nrows", CASES_PERTURBED, ids=[c[0] for c in CASES_PERTURBED])
def test_perturbed_runs_through_the_shim(name, exe, args, ours_f, brute_f, tol_f, nrows, suffix, prefix, tmp_path):
    run_test(exe + "_cuda" + suffix, args, tmp_path, timeout=90)
    ours = load_table(tmp_path / ours_f)
    brute = load_table(tmp_path / brute_f)
    tol = load_table(os.path.join(REF, tol_f))
    bad = difffloat(ours, brute, tol, nrows)     # the reference's gate: the main cell's n^3 rows (all for the periodic test)
    assert len(bad) == 0, "difffloat gate: %d entries out of tolerance, first (row, col) %s" % (len(bad), bad[:5])
    ref = np.load(os.path.join(GOLD, prefix + name + ".npz"))["table"]
    bad = difffloat(ours, ref, tol, 10 ** 6)
    assert len(bad) == 0, "against the reference's own table: %d entries, first (row, col) %s" % (len(bad), bad[:5])
