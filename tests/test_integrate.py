"""Time integration either side of the neighbour loops: drift, kick1/kick2, time-step.

CPU (-m "not gpu"): the C restatement (oracle_drift / oracle_kick / oracle_timestep) against
  * the golden states in tests/golden/integ/, written by the REFERENCE's own compiled cell_drift_part,
    runner_do_kick1/2, runner_do_timestep and hydro operators (tests/golden/make_golden_integ.py), and
  * the compiled reference live (oracle/_ref), per operator on adversarial states and over engine-step loops.
GPU (-m gpu): the CUDA kernels through the C ABI against the oracle and the golden states, every field after
every step, bit for bit: the particles go to the device once and stay there."""
import os

import numpy as np
import pytest

import integ_lib as il
import oracle_lib as ol

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "integ")


def _case(name):
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_integ",
                                                  os.path.join(os.path.dirname(GOLD), "make_golden_integ.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.case(name)


CASES = ["clustered4k_bins", "sedov12_big_steps", "impact3k_planetary"]


def _assert_state_equal(a, b, what, keys=il.STATE_KEYS):
    for k in keys:
        assert np.array_equal(np.asarray(a[k]).reshape(-1), np.asarray(b[k]).reshape(-1)), \
            "%s: field %s is not bit-identical" % (what, k)


# --------------------------------------------------------------------------- CPU
def test_timeline_known_answers():
    """src/timeline.h: get_max_active_bin, and the bins of a first step (tests/testTimeIntegration-style
    arithmetic: a step of bin b is 2^(b+1) ticks)."""
    assert il.get_max_active_bin(0) == il.NUM_TIME_BINS
    assert il.get_max_active_bin(1 << 44) == 43 and il.get_max_active_bin(3 << 44) == 43
    assert il.get_max_active_bin(1 << 45) == 44
    n = 4
    ic = {"x": np.zeros((n, 3)), "v": np.zeros((n, 3), np.float32), "u": np.ones(n, np.float32),
          "h": np.full(n, 0.1, np.float32)}
    s = il.new_state(ic)
    s["rho"][:] = 1.0
    s["v_sig"][:] = np.array([1.0, 2.0, 4.0, 1e-3], dtype=np.float32)
    ip = il.default_iparams(cfl=0.1, dt_max=1e-2)
    chk = il.Checker("oracle", ip)
    t_end, updated = chk.timestep(s, 0, il.NUM_TIME_BINS)
    assert updated == n
    # dt_cfl = 2 * 1.825742 * 0.1 * 0.1 / v_sig = 0.0365 / v_sig, capped by dt_max; bin = floor(log2(dt / time_base)) - 1
    want = []
    for vs in (1.0, 2.0, 4.0, 1e-3):
        dt = min(2 * 1.825742 * 0.1 * 0.1 / vs, 1e-2)
        want.append(int(np.floor(np.log2(dt / ip["time_base"]))) - 1)
    assert list(s["time_bin"]) == want
    assert t_end == 1 << (min(want) + 1)


@pytest.mark.parametrize("name", CASES)
def test_oracle_loop_reproduces_reference_golden(name):
    icc, tree, kw, ip, variant, nsteps, eos = _case(name)
    z = np.load(os.path.join(GOLD, name + ".npz"))
    seen = {"expf": 0}

    def watch(step, s):
        seen["h_dt"], seen["h"] = s["h_dt"].copy(), s["h"].copy()

    s, log = il.run_loop("oracle", icc, kw, ip, tree, nsteps, eos=eos, variant=variant, on_step=watch)
    _assert_state_equal(s, z, name)
    assert [[l[0], l[1], l[2], l[5]] for l in log] == z["log_int"].tolist()
    assert np.array_equal(np.array(log, dtype=np.float64), z["log"]), "dx_max differ"
    if name != "impact3k_planetary":
        assert len(np.unique(s["time_bin"])) >= 4, "not a time-bin hierarchy"


def test_golden_sedov_case_takes_the_expf_branch():
    """hydro_predict_extra switches from the 4th-order series to libm's expf when |h_dt / h * dt| >= 0.2; the
    big-step fixture must exercise it (the CUDA side carries its own glibc-exact expf for that branch)."""
    icc, tree, kw, ip, variant, nsteps, eos = _case("sedov12_big_steps")
    hits = []

    def watch(step, s):
        hits.append((s["h_dt"].copy(), s["h"].copy()))

    s, log = il.run_loop("oracle", icc, kw, ip, tree, nsteps, on_step=watch)
    n_big = 0
    for k in range(len(log) - 1):
        dt = np.float32((log[k + 1][0] - log[k][0]) * ip["time_base"])
        n_big += int((np.abs(hits[k][0] / hits[k][1] * dt) >= 0.2).sum())
    assert n_big > 0


def _adversarial_state(n, seed, planetary):
    rng = np.random.default_rng(seed)
    ic = {"x": rng.random((n, 3)), "v": rng.standard_normal((n, 3)).astype(np.float32),
          "u": (rng.random(n) * 2 + 0.01).astype(np.float32), "h": (rng.random(n) * 0.2 + 0.01).astype(np.float32)}
    if planetary:
        ic["mat_id"] = rng.choice(np.array([0, 100, 101], dtype=np.int32), n)
        ic["u"] = (rng.random(n) * 5 + 0.1).astype(np.float32)
    s = il.new_state(ic)
    s["rho"] = (rng.random(n) * (8.0 if planetary else 2.0) + 0.5).astype(np.float32)
    s["v_sig"] = (rng.random(n) * 3 + 0.1).astype(np.float32)
    s["a"] = (rng.standard_normal((n, 3)) * 10).astype(np.float32)
    s["u_dt"] = (rng.standard_normal(n) * 50).astype(np.float32)      # some hit the "no more than halved" limit
    s["h_dt"] = (rng.standard_normal(n) * 3 * s["h"]).astype(np.float32)  # |w1| beyond 0.2 for some: expf branch
    s["h_dt"][::7] = 0.0                                               # dt_h_change = FLT_MAX
    s["v_full"] = (s["v"] + rng.standard_normal((n, 3)).astype(np.float32) * 0.1).astype(np.float32)
    s["u_full"] = (s["u"] * (1 + 0.1 * rng.standard_normal(n))).astype(np.float32)
    s["x_diff"] = (rng.standard_normal((n, 3)) * 1e-3).astype(np.float32)
    s["x_diff_sort"] = (rng.standard_normal((n, 3)) * 1e-3).astype(np.float32)
    s["time_bin"] = rng.integers(44, 50, n).astype(np.int8)
    s["min_ngb_time_bin"] = rng.integers(42, 58, n).astype(np.int8)
    return s


@pytest.mark.parametrize("variant", ["minimal", "planetary"])
def test_operators_oracle_vs_compiled_reference(variant):
    """Each operator on its own, states built to reach every branch: series and expf in the drift, the h clamp,
    the energy floor and the halving limit in the kick, neighbour-limited, doubled and dt_max-capped steps."""
    if not ol.ref_available(variant):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    planetary = variant == "planetary"
    eos = ol.eos_from_json() if planetary else None
    ip = il.default_iparams(scheme=int(planetary), dt_max=3e-2, h_max=0.18, h_min=0.02, min_u=0.05, time_end=1024.0)
    chk_o = il.Checker("oracle", ip, eos=eos, variant=variant)
    chk_r = il.Checker("ref", ip, eos=eos, variant=variant)
    ti_old, ti_current = 5 << 46, 6 << 46
    mab = il.get_max_active_bin(ti_current)
    for op in ("drift", "kick1", "kick2", "timestep"):
        so = _adversarial_state(3000, 17, planetary)
        sr = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in so.items()}
        if op == "drift":
            ro, rr = chk_o.drift(so, ti_old, ti_current), chk_r.drift(sr, ti_old, ti_current)
            w1 = np.abs(sr["h_dt"] / sr["h"]) * (ti_current - ti_old) * ip["time_base"]
            assert (w1 >= 0.2).any() and (w1 < 0.2).any()
        elif op == "timestep":
            ro, rr = chk_o.timestep(so, ti_current, mab), chk_r.timestep(sr, ti_current, mab)
        else:
            which = 1 if op == "kick1" else 2
            ro, rr = chk_o.kick(so, which, ti_current, mab), chk_r.kick(sr, which, ti_current, mab)
        assert ro == rr, op
        _assert_state_equal(so, sr, op)
        if op != "drift":
            act = so["time_bin"] <= mab
            assert act.any() and not act.all()


@pytest.mark.parametrize("name", CASES[:1])
def test_oracle_loop_vs_compiled_reference_live(name):
    if not ol.ref_available("minimal"):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    icc, tree, kw, ip, variant, nsteps, eos = _case(name)
    so, lo = il.run_loop("oracle", icc, kw, ip, tree, 3, eos=eos, variant=variant)
    sr, lr = il.run_loop("ref", icc, kw, ip, tree, 3, eos=eos, variant=variant)
    assert lo == lr
    _assert_state_equal(so, sr, name)


# --------------------------------------------------------------------------- GPU
def _gpu_backend(kw, eos):
    import parity
    from swift_subtask_dev_b200 import HydroBackend
    k = dict(kw)
    k["max_active_bin"] = il.NUM_TIME_BINS
    return HydroBackend(parity.gpu_params(k, eos))


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_resident_loop_matches_oracle_and_golden(name):
    """drift / kick / timestep kernels + the neighbour loops over several engine steps with individual time
    bins, the state compared with the oracle's after EVERY step and with the reference's golden state at the
    end: x, v, v_full, u, u_full, h, rho, P, c, v_sig, a, u_dt, h_dt, x_diff, x_diff_sort, time_bin."""
    icc, tree, kw, ip, variant, nsteps, eos = _case(name)
    ref_states = []
    so, lo = il.run_loop("oracle", icc, kw, ip, tree, nsteps, eos=eos, variant=variant,
                         on_step=lambda step, s: ref_states.append({k: s[k].copy() for k in il.STATE_KEYS}))
    be = _gpu_backend(kw, eos)

    def check(step, g):
        _assert_state_equal(g, ref_states[step], "%s step %d" % (name, step))

    sg, lg = il.run_loop_gpu(be, icc, ip, tree, nsteps, on_step=check)
    t = be.integrator_timings()
    assert all(t[k] > 0.0 for k in ("drift_ms", "kick1_ms", "kick2_ms", "timestep_ms"))
    # after a re-sort the host zeroes x_diff_sort only (runner_do_hydro_sort: xp->x_diff_sort = 0), after a
    # rebuild both (space_rebuild)
    assert np.abs(sg["x_diff_sort"]).max() > 0
    be.reset_offsets(1)
    s1 = be.state()
    assert not s1["x_diff_sort"].any() and np.array_equal(s1["x_diff"], sg["x_diff"])
    be.reset_offsets(2)
    assert not be.state()["x_diff"].any()
    be.close()
    assert [l[:3] + l[5:] for l in lg] == [l[:3] + l[5:] for l in lo], "active counts / sync points differ"
    assert np.array_equal(np.array(lg, dtype=np.float32), np.array(lo, dtype=np.float32)), "dx_max differ"
    z = np.load(os.path.join(GOLD, name + ".npz"))
    _assert_state_equal(sg, z, name + " vs reference golden")


@pytest.mark.gpu
def test_gpu_rebuild_and_reinit_reproduce_the_run():
    """A rebuild frees the integrator's device state with everything else; building and initialising again on
    the same context (what a host does after space_rebuild) gives the same run, bit for bit."""
    icc, tree, kw, ip, variant, nsteps, eos = _case("clustered4k_bins")
    be = _gpu_backend(kw, eos)
    s1, l1 = il.run_loop_gpu(be, icc, ip, tree, 3)
    s2, l2 = il.run_loop_gpu(be, icc, ip, tree, 3)     # build_from_cells + integrator_init again
    be.integrator_init(**ip)                            # a second init on the same build is allowed
    be.close()
    assert l1 == l2
    _assert_state_equal(s1, s2, "second run on the same context")


@pytest.mark.gpu
def test_gpu_integrator_needs_all_cells_in_the_work_lists():
    """A context built with part of the particles inactive has dropped their cells from the work lists."""
    import parity
    from swift_subtask_dev_b200 import HydroBackend, ics
    from swift_subtask_dev_b200.backend import SwiftCudaError
    ic = ics.uniform_random(20 ** 3, seed=2)
    ic["time_bin"] = np.where(ic["x"][:, 0] < 0.5, 1, 5).astype(np.int8)
    be = HydroBackend(parity.gpu_params(dict(periodic=True, max_active_bin=2)))
    be.build(ic)
    with pytest.raises(SwiftCudaError):
        be.integrator_init(**il.default_iparams())
    be.close()
