"""swift_cuda_timestep_static (include/swift_cuda.h): the allocation-free variant of the time-step kernel,
its reduction scratch in a file-scope device array of the library, against the existing entry
swift_cuda_timestep on the same particles.  Integer work: new time bins, the next sync point and the
number of updated particles must be IDENTICAL, on two calls in a row (the second call finds the scratch
as the first left it), over two engine steps with individual time bins.

Collected last on purpose: passes alone in a fresh process and after
the other GPU tests."""
import numpy as np
import pytest

import integ_lib as il
from swift_subtask_dev_b200 import HydroBackend, ics, make_params
from swift_subtask_dev_b200 import backend as B


def _integrator(ip):
    return B.Integrator(ip["cfl_condition"], ip["log_max_h_change"], ip["minimal_internal_energy"], ip["dt_min"],
                        ip["dt_max"], ip["time_base"])


def _static_timestep(torch, st, ti_current, mab, integ, stream=None):
    """Uploads the inputs of the time-step (caller order, as state() gives them) into tensors the TEST owns,
    runs the allocation-free entry on them and returns (ti_end_min, updated, new time bins)."""
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(np.ascontiguousarray(st[k])).to(dev) for k in
         ("h", "v_sig", "h_dt", "min_ngb_time_bin", "time_bin")}
    torch.cuda.synchronize()
    n = int(t["h"].numel())
    before = torch.cuda.memory_allocated(dev)
    ti_end_min, updated = B.timestep_static(0, n, t["h"].data_ptr(), t["v_sig"].data_ptr(), t["h_dt"].data_ptr(),
                                            t["min_ngb_time_bin"].data_ptr(), t["time_bin"].data_ptr(), ti_current,
                                            mab, integ, stream)
    assert torch.cuda.memory_allocated(dev) == before
    return ti_end_min, updated, t["time_bin"].cpu().numpy()


@pytest.mark.gpu
def test_timestep_static_matches_timestep_on_two_calls_in_a_row():
    import torch
    ic = ics.clustered(40000, nclumps=12, seed=21)
    ic["time_bin"] = np.zeros(len(ic["h"]), dtype=np.int8)
    ip = il.default_iparams(dt_max=1e-2, time_end=1.0)
    integ = _integrator(ip)
    be = HydroBackend(make_params(periodic=True, max_active_bin=il.NUM_TIME_BINS))
    be.build(ic)
    be.integrator_init(**ip)
    side = torch.cuda.Stream(device=0)
    ti_current, ti_old = 0, 0
    calls = 0
    for step in range(3):
        mab = il.get_max_active_bin(ti_current)
        be.set_time(ti_current, mab)
        if step > 0:
            be.drift(ti_old)
        be.step(do_sort=True)
        if step > 0:
            be.kick(2)
        before = be.state()
        want_t, want_u = be.timestep()           # the existing entry
        after = be.state()
        if step > 0:
            assert 0 < want_u < len(ic["h"]), "the case should have active and inactive particles"
        # the new entry, twice in a row on fresh copies of the same inputs: default stream, then a side stream
        for stream in (None, side.cuda_stream):
            got_t, got_u, bins = _static_timestep(torch, before, ti_current, mab, integ, stream)
            calls += 1
            assert (got_t, got_u) == (want_t, want_u), (step, calls, got_t, got_u, want_t, want_u)
            assert np.array_equal(bins, after["time_bin"]), "step %d call %d: time bins differ" % (step, calls)
        be.kick(1)
        ti_old, ti_current = ti_current, want_t
    be.close()
    assert calls == 6


@pytest.mark.gpu
def test_timestep_static_rejects_what_it_cannot_hold():
    import torch
    integ = _integrator(il.default_iparams())
    x = torch.zeros(8, dtype=torch.float32, device="cuda:0")
    b = torch.zeros(8, dtype=torch.int8, device="cuda:0")
    with pytest.raises(B.SwiftCudaError, match="exceeds the 16777216 particles"):
        B.timestep_static(0, B.TIMESTEP_STATIC_MAX + 1, x.data_ptr(), x.data_ptr(), x.data_ptr(), b.data_ptr(),
                          b.data_ptr(), 0, 56, integ)
    # a host pointer where a device pointer belongs is refused, not dereferenced
    host = np.zeros(8, dtype=np.float32)
    with pytest.raises(B.SwiftCudaError, match="not a device pointer"):
        B.timestep_static(0, 8, host.ctypes.data, x.data_ptr(), x.data_ptr(), b.data_ptr(), b.data_ptr(), 0, 56, integ)
    # an empty call is fine and touches nothing
    assert B.timestep_static(0, 0, None, None, None, None, None, 0, 56, integ) == (1 << 57, 0)
