"""The time-step limiter (swift --limiter; the remainder of SURVEY.md 8(f) N3), CPU side: the plain-C restatement
(oracle/sph_oracle.c: stage 4 of the density-shaped walks, oracle_limiter) against

  * the golden fixtures in tests/golden/limiter/, written by the REFERENCE's own compiled
    runner_do{self,pair}1_branch_limiter / runner_dosub_*1_limiter and runner_do_limiter
    (tests/golden/make_golden_limiter.py), and
  * the compiled reference itself, live, where oracle/_ref is present (the build container),

plus known answers of timestep_limit_part worked out by hand from src/timestep_limiter.h:67-217.
The GPU tests of the same fixtures are in tests/test_zyz_gpu_limiter.py."""
import os

import numpy as np
import pytest

import integ_lib as il
import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "limiter")


def gen():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_limiter",
                                                  os.path.join(os.path.dirname(GOLD), "make_golden_limiter.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


LOOP_CASES = ["loop_clustered3k_open", "loop_uniform12_periodic", "loop_lattice10_all_starting"]
RUN_CASES = ["run_sedov14", "run_sedov12_sphenix"]


def brute_force_wakeup(x, h, bins, mab, dim, periodic, same_leaf=None):
    """What the limiter loop computes, from its definition: every starting i wakes every j within its support
    whose bin is more than 2 above -- except, for two particles of the same leaf cell (DOSELF1), when j starts
    too and i lies in ITS support (the symmetric runner_iact_limiter does nothing)."""
    n = len(h)
    wake = np.full(n, il.NOT_AWAKE, dtype=np.int64)
    gamma = np.float32(1.825742)
    hg2 = (h * h * np.float32(gamma * gamma)).astype(np.float32)
    for i in np.nonzero(bins <= mab)[0]:
        dx = x[i] - x
        if periodic:
            dx -= np.round(dx / dim) * dim
        r2 = (dx.astype(np.float32) ** 2).sum(1)
        hit = (r2 < hg2[i]) & (bins > bins[i] + 2)
        hit[i] = False
        if same_leaf is not None:
            hit &= ~(same_leaf[i] & (bins <= mab) & (r2 < hg2))
        wake[hit] = np.maximum(wake[hit], -int(bins[i]))
    return wake


@pytest.mark.parametrize("name", LOOP_CASES)
def test_oracle_limiter_loop_reproduces_reference_golden(name):
    m = gen()
    icc, tree, kw = m.loop_case(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    # the fixture belongs to these inputs
    assert np.array_equal(z["time_bin"], icc["time_bin"]) and np.array_equal(z["h"], icc["h"])
    assert np.array_equal(z["x"], icc["x"])
    w = m.run_loop_only("oracle", icc, tree, kw)
    assert np.array_equal(w, z["l_wakeup"])
    woken = w != il.NOT_AWAKE
    assert woken.sum() > len(w) // 4
    # nobody is woken by somebody slower, and a woken particle was more than 2 bins above its caller
    assert (icc["time_bin"][woken].astype(int) > -w[woken].astype(int) + 2).all()


@pytest.mark.parametrize("name", LOOP_CASES[:2])
def test_limiter_loop_against_brute_force(name):
    """All pairs, no tree, no sorts: the wake-up flags follow from positions, h and bins alone.  (In these two
    cases no starting particle can be woken -- max_active_bin + 1 <= bin_i + 2 fails for none of them -- so the
    DOSELF1 exception does not enter.)"""
    m = gen()
    icc, tree, kw = m.loop_case(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    want = brute_force_wakeup(icc["x"], icc["h"], icc["time_bin"].astype(np.int64), kw["max_active_bin"], 1.0,
                              kw["periodic"])
    assert np.array_equal(want, z["l_wakeup"].astype(np.int64))


def test_limiter_loop_same_cell_exception():
    """Two starting particles that see each other: in ONE cell (DOSELF1) nothing happens, in a PAIR of cells the
    slower one is woken (src/runner_doiact_functions_limiter.h:655-660 against :353-356, 437-440)."""
    prm_kw = dict(periodic=False, max_active_bin=9)
    # h = 0.05: a 9x9x9 top grid (cell width 1/9 > 1.1 gamma h), support gamma h = 0.091; two particles 0.02 apart
    h = np.float32(0.05)

    def run(xa, xb):
        x = np.array([xa, xb], dtype=np.float64)
        ic = {"x": x, "v": np.zeros((2, 3), np.float32), "mass": np.ones(2, np.float32),
              "h": np.full(2, h, np.float32), "u": np.ones(2, np.float32), "id": np.arange(2, dtype=np.int64)}
        prm = ol.make_params(**prm_kw)
        order, cells, cdim, tasks, nt = ol.build_tree_and_tasks(prm, ic["x"], ic["h"])
        from swift_subtask_dev_b200 import ics
        icc = ics.permute(ic, order)
        bins = np.array([1, 6], dtype=np.int8)[order]
        icc["time_bin"] = bins
        icc["l_time_bin"], icc["l_only"] = bins, True
        out = {}
        for which in ("oracle",) + (("ref",) if ol.ref_available("minimal") else ()):
            w = ol.run_step(which, prm, icc, cells, tasks, nt)["l_wakeup"]
            back = np.empty(2, dtype=np.int8)
            back[order] = w
            out[which] = back
        for w in out.values():
            assert np.array_equal(w, out["oracle"])
        return cdim, out["oracle"]

    cdim, same = run([0.49, 0.5, 0.5], [0.51, 0.5, 0.5])        # both in the cell [4/9, 5/9)
    assert tuple(cdim) == (9, 9, 9)
    assert list(same) == [il.NOT_AWAKE, il.NOT_AWAKE]
    cdim, across = run([0.545, 0.5, 0.5], [0.565, 0.5, 0.5])    # the cell boundary is at 5/9
    assert list(across) == [il.NOT_AWAKE, -1]
    # and when the slower one does not start a step (bin 6 > max_active_bin) it is woken in one cell too
    prm_kw["max_active_bin"] = 3
    cdim, same = run([0.49, 0.5, 0.5], [0.51, 0.5, 0.5])
    assert list(same) == [il.NOT_AWAKE, -1]


def _one_particle_state(time_bin, v_full, a, u_full, u_dt):
    s = il.new_state({"x": np.zeros((1, 3)), "v": np.zeros((1, 3), np.float32), "u": np.ones(1, np.float32),
                      "h": np.ones(1, np.float32)})
    s["time_bin"][:] = time_bin
    s["v_full"][0] = v_full
    s["a"][0] = a
    s["u_full"][0] = u_full
    s["u_dt"][0] = u_dt
    return s


def test_timestep_limit_part_known_answers():
    """src/timestep_limiter.h:67-217 by hand.  Time line: time_base = 2^-20; a particle in bin 10 (step 2^11 =
    2048 ticks) started at tick 0 and has had its kick1 over 1024 ticks; at ti_current = 512 (max_active_bin 8)
    a neighbour of bin 2 wakes it: new bin 4 (step 32)."""
    tb = 2.0 ** -20
    ip = dict(il.default_iparams(), time_base=tb)
    for which in ("oracle",) + (("ref",) if ol.ref_available("minimal") else ()):
        chk = il.Checker(which, ip)
        # inactive, new bin 4 <= max_active_bin 8: kick1 undone (-1024 ticks), +512 ticks to ti_beg_new = 512
        # (512 is a multiple of 32), no new kick1: the kick task does it.  Net: -512 ticks of a and u_dt.
        s = _one_particle_state(10, [1.0, 0.0, 0.0], [2.0, 0.0, -4.0], 8.0, 16.0)
        wake = np.array([-2], dtype=np.int8)
        t_end, nlim = chk.limiter(s, wake, 512, 8)
        assert nlim == 1 and wake[0] == il.NOT_AWAKE and s["time_bin"][0] == 4
        assert t_end == 512 + 32
        dt = 512 * tb
        assert s["v_full"][0].tolist() == [1.0 - 2.0 * dt, 0.0, 4.0 * dt]
        assert s["u_full"][0] == np.float32(8.0 - 16.0 * dt)
        # the same at ti_current = 520 (max_active_bin 2): ti_beg_new = 512, the bin-4 step is under way, so its
        # kick1 (16 ticks) is applied on the spot: net -1024 + 512 + 16 ticks; the step ends at 544
        s = _one_particle_state(10, [1.0, 0.0, 0.0], [2.0, 0.0, -4.0], 8.0, 16.0)
        wake = np.array([-2], dtype=np.int8)
        t_end, nlim = chk.limiter(s, wake, 520, 2)
        assert s["time_bin"][0] == 4 and t_end == 544
        dt = (1024 - 512 - 16) * tb
        assert np.allclose(s["v_full"][0], [1.0 - 2.0 * dt, 0.0, 4.0 * dt], rtol=0, atol=1e-7)
        assert abs(float(s["u_full"][0]) - (8.0 - 16.0 * dt)) < 1e-6
        # an ACTIVE particle (bin 8 at ti_current = 512) only changes the length of the step that starts
        s = _one_particle_state(8, [1.0, 0.0, 0.0], [2.0, 0.0, -4.0], 8.0, 16.0)
        wake = np.array([-3], dtype=np.int8)
        t_end, nlim = chk.limiter(s, wake, 512, 8)
        assert s["time_bin"][0] == 5 and t_end == 512 + 64
        assert s["v_full"][0].tolist() == [1.0, 0.0, 0.0] and s["u_full"][0] == 8.0
        # nobody woken: nothing happens, the sentinel comes back
        s = _one_particle_state(10, [1.0, 0.0, 0.0], [2.0, 0.0, -4.0], 8.0, 16.0)
        wake = np.array([il.NOT_AWAKE], dtype=np.int8)
        t_end, nlim = chk.limiter(s, wake, 512, 8)
        assert nlim == 0 and t_end == 1 << (il.NUM_TIME_BINS + 1) and s["time_bin"][0] == 10


def test_oracle_limiter_reproduces_reference_golden():
    z = np.load(os.path.join(GOLD, "apply_sedov14.npz"))
    ti_current, mab, t_end, nlim = (int(v) for v in z["scalars"])
    m = gen()
    ip = m.RUN_CASES["run_sedov14"][2]
    s = {k: z["before_" + k].copy() for k in il.STATE_KEYS}
    s["min_ngb_time_bin"] = np.full(len(s["h"]), il.NUM_TIME_BINS + 1, dtype=np.int8)
    s["mat_id"] = None
    wake = z["wake"].copy()
    assert il.limiter_census(wake, s["time_bin"], mab) == tuple(int(v) for v in z["census"])
    assert min(z["census"]) > 100, "every branch of timestep_limit_part in quantity"
    got = il.Checker("oracle", ip).limiter(s, wake, ti_current, mab)
    assert got == (t_end, nlim)
    assert (wake == il.NOT_AWAKE).all()
    for k in il.STATE_KEYS:
        assert np.array_equal(s[k], z["after_" + k]), k
    changed = [k for k in il.STATE_KEYS if not np.array_equal(z["before_" + k], z["after_" + k])]
    assert sorted(changed) == ["time_bin", "u_full", "v_full"]


@pytest.mark.parametrize("name", RUN_CASES)
def test_oracle_run_with_limiter_reproduces_reference_golden(name):
    m = gen()
    icc, tree, kw, ip, variant, nsteps = m.run_case(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    sx = ol.SPHENIX_PROPS if variant == "sphenix" else None
    s, log = il.run_loop("oracle", icc, kw, ip, tree, nsteps, variant=variant, sphenix=sx, limiter=True)
    for k in il.STATE_KEYS:
        assert np.array_equal(np.asarray(s[k]).reshape(-1), z[k].reshape(-1)), k
    assert [[l[0], l[1], l[2], l[5], l[6], l[7], l[8]] for l in log] == z["log_int"].tolist()
    assert np.array_equal(np.array([l[:6] for l in log], dtype=np.float64), z["log"])
    woken = z["log_int"][:, 4:].sum(0)
    assert woken[0] > 0 and woken[1] > 0 and woken[2] > 0
    # the limiter changes the run: without it the same steps end elsewhere
    s0, log0 = il.run_loop("oracle", icc, kw, ip, tree, nsteps, variant=variant, sphenix=sx)
    assert not np.array_equal(s0["time_bin"], s["time_bin"])


@pytest.mark.skipif(not ol.ref_available("minimal"), reason="oracle/_ref not built (needs the reference tree)")
def test_oracle_run_with_limiter_vs_compiled_reference_live():
    """A case no fixture holds: clustered particles (deep tree, sub tasks), energy spike in the densest clump."""
    from swift_subtask_dev_b200 import ics
    ic = ics.clustered(4000, nclumps=4, seed=17)
    ic["u"] = ic["u"].copy()
    ic["u"][np.argsort(ic["h"])[:10]] *= 3e3
    kw = dict(periodic=True)
    icc, tree = il.prepare(ic, kw)
    ip = il.default_iparams()
    sr, lr = il.run_loop("ref", icc, kw, ip, tree, 8, limiter=True)
    so, lo = il.run_loop("oracle", icc, kw, ip, tree, 8, limiter=True)
    assert lr == lo
    assert sum(l[6] for l in lr) > 0, "nobody was woken: the case does not exercise the limiter"
    for k in il.STATE_KEYS:
        assert np.array_equal(sr[k], so[k]), k


# ---- the body of the device kernel k_limiter, compiled for the host ---------------------------------------
@pytest.fixture(scope="module")
def kernel_body(tmp_path_factory):
    """swift_subtask_dev_b200/csrc/limit_part.h -- the source nvcc compiles into k_limiter -- built by g++ without
    contraction behind tests/host_kernels/limit_part_host.cpp."""
    import ctypes as C
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    so = str(tmp_path_factory.mktemp("host_kernels") / "liblimitpart.so")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-o", so,
                    os.path.join(here, "host_kernels", "limit_part_host.cpp")], check=True)
    lib = C.CDLL(so)
    lib.limit_parts_host.argtypes = [C.c_int, C.c_longlong, C.c_double, C.c_int, C.c_float] + [C.c_void_p] * 8
    lib.limit_parts_host.restype = C.c_int

    def run(s, wake, ti_current, mab, ip):
        import ctypes as C
        t, u = C.c_longlong(0), C.c_longlong(0)
        energy_min = max(np.float32(ip["minimal_internal_energy"]), np.float32(0.0))
        rc = lib.limit_parts_host(len(s["h"]), ti_current, ip["time_base"], mab, float(energy_min), wake.ctypes.data,
                                  s["time_bin"].ctypes.data, s["v_full"].ctypes.data, s["u_full"].ctypes.data,
                                  s["a"].ctypes.data, s["u_dt"].ctypes.data, C.addressof(t), C.addressof(u))
        assert rc == 0
        return t.value, u.value
    return run


def test_kernel_body_on_the_host_reproduces_reference_golden(kernel_body):
    z = np.load(os.path.join(GOLD, "apply_sedov14.npz"))
    ti_current, mab, t_end, nlim = (int(v) for v in z["scalars"])
    ip = gen().RUN_CASES["run_sedov14"][2]
    s = {k: np.ascontiguousarray(z["before_" + k]).copy() for k in il.STATE_KEYS}
    wake = z["wake"].copy()
    assert kernel_body(s, wake, ti_current, mab, ip) == (t_end, nlim)
    assert (wake == il.NOT_AWAKE).all()
    for k in il.STATE_KEYS:
        assert np.array_equal(s[k], z["after_" + k]), k


def test_kernel_body_on_the_host_vs_oracle_at_the_energy_floor(kernel_body):
    """Strongly cooling particles: undoing a kick cannot take u_full below half (hydro_kick_extra), and with a
    minimal internal energy u_dt is zeroed on the way -- the later kicks of the same call then see u_dt = 0."""
    z = np.load(os.path.join(GOLD, "apply_sedov14.npz"))
    ti_current, mab = int(z["scalars"][0]), int(z["scalars"][1])
    ip = dict(gen().RUN_CASES["run_sedov14"][2])
    rng = np.random.default_rng(3)
    base = {k: np.ascontiguousarray(z["before_" + k]).copy() for k in il.STATE_KEYS}
    n = len(base["h"])
    dt_old = (1 << (base["time_bin"].astype(np.int64) + 1)) * ip["time_base"]
    # |u_dt| * dt comparable with u_full, either sign
    base["u_dt"] = (rng.uniform(-4.0, 4.0, n) * base["u_full"] / dt_old).astype(np.float32)
    ip["minimal_internal_energy"] = float(np.median(base["u_full"]))
    a = {k: v.copy() for k, v in base.items()}
    b = {k: v.copy() for k, v in base.items()}
    b["min_ngb_time_bin"] = np.full(n, il.NUM_TIME_BINS + 1, dtype=np.int8)
    b["mat_id"] = None
    wa, wb = z["wake"].copy(), z["wake"].copy()
    ra = kernel_body(a, wa, ti_current, mab, ip)
    rb = il.Checker("oracle", ip).limiter(b, wb, ti_current, mab)
    assert ra == rb
    for k in il.STATE_KEYS:
        assert np.array_equal(a[k], b[k]), k
    assert (a["u_dt"] == 0).sum() > (base["u_dt"] == 0).sum() + 50, "the floor was not reached"
    if ol.ref_available("minimal"):
        c = {k: v.copy() for k, v in base.items()}
        c["min_ngb_time_bin"], c["mat_id"] = b["min_ngb_time_bin"], None
        wc = z["wake"].copy()
        assert il.Checker("ref", ip).limiter(c, wc, ti_current, mab) == ra
        for k in il.STATE_KEYS:
            assert np.array_equal(a[k], c[k]), k + " (compiled reference)"


@pytest.mark.skipif(not ol.ref_available("planetary"), reason="oracle/_ref not built (needs the reference tree)")
def test_oracle_run_with_limiter_planetary_vs_compiled_reference_live():
    """The Planetary scheme (Tillotson): kick_part's energy floor is the scheme's own (hydro_kick_extra,
    src/hydro/Planetary/hydro.h:904-925), and the undone / redone kicks of a woken particle go through it."""
    from swift_subtask_dev_b200 import ics
    ic = ics.two_body_impact(3000, seed=5)
    kw = dict(periodic=False, dim=tuple(float(v) for v in ic["dim"]), scheme=1, alpha=1.5, h_max=float(ic["h_max"]))
    ip = il.default_iparams(scheme=1, h_max=float(ic["h_max"]), dt_max=1e3, time_end=1e5)
    # a hot spot on the impactor's surface: its neighbours fall more than two bins behind
    hot = np.argsort(ic["x"][:, 0])[-12:]
    ic["u"] = ic["u"].copy()
    ic["u"][hot] *= 400.0
    eos = ol.eos_from_json()
    icc, tree = il.prepare(ic, kw)
    sr, lr = il.run_loop("ref", icc, kw, ip, tree, 6, eos=eos, variant="planetary", limiter=True)
    so, lo = il.run_loop("oracle", icc, kw, ip, tree, 6, eos=eos, variant="planetary", limiter=True)
    assert lr == lo
    assert sum(l[6] for l in lr) > 0, "nobody was woken: the case does not exercise the limiter"
    for k in il.STATE_KEYS:
        assert np.array_equal(sr[k], so[k]), k
