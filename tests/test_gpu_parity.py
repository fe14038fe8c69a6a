"""GPU parity tests: the CUDA path (through the C ABI) against the oracle on
the same seeded inputs.  Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

import oracle_lib as ol
import parity
from swift_subtask_dev_b200 import ics

pytestmark = pytest.mark.gpu

BASE = dict(periodic=True)


def _check(ic, kw, pairs=True, n_boundary_flips_allowed=None):
    cap = len(ic["h"]) * 200 if pairs else 0
    g = parity.run_gpu(ic, kw, record=pairs, pair_capacity=cap)
    o = parity.run_oracle_on_tree(ic, g["tree"], kw, record=pairs)
    r = parity.compare(g, o)
    # first density pass: identical inputs -> neighbour counts and pair lists bit-exact
    assert r["count_density"][0] == r["count_density"][1]
    if pairs:
        assert parity.pair_set_diff(g["pairs"][0], o["pairs"][0]) == 0
    for k in ("d_rho", "d_wcount"):
        assert r[k] <= parity.TOL_RHO_H, (k, r[k])
    for k in ("d_rho_dh", "d_wcount_dh", "d_div_v", "d_rot_v"):
        assert r[k] <= parity.TOL_FORCE, (k, r[k])
    assert r["h"] <= parity.TOL_RHO_H and r["rho"] <= parity.TOL_RHO_H
    for k in ("pressure", "soundspeed", "f", "balsara", "a", "u_dt", "h_dt", "v_sig"):
        assert r[k] <= parity.TOL_FORCE, (k, r[k])
    # Later loops see h values that went through differently ordered f32 sums, so a pair
    # sitting within an ulp of r2 == h^2 gamma^2 may flip; anything else must agree.
    n = len(ic["h"])
    allowed = max(4, n // 20000) if n_boundary_flips_allowed is None else n_boundary_flips_allowed
    assert abs(r["count_redo"][0] - r["count_redo"][1]) <= allowed
    assert abs(r["count_force"][0] - r["count_force"][1]) <= allowed
    if pairs:
        assert parity.pair_set_diff(g["pairs"][2], o["pairs"][2]) <= 2 * allowed
    return g, o, r


def test_lattice_16():
    _check(ics.lattice(16), dict(BASE))


def test_uniform_random_32():
    _check(ics.uniform_random(32 ** 3, seed=11), dict(BASE))


def test_uniform_random_split_cells():
    # 80^3 in 12^3 top cells: ~296 per cell... use 96^3 -> 512 per top cell -> split once
    _check(ics.uniform_random(96 ** 3, seed=5), dict(BASE), pairs=False)


def test_clustered_deep_tree():
    _check(ics.clustered(60000, nclumps=16, seed=3), dict(BASE))


def test_non_periodic():
    _check(ics.uniform_random(28 ** 3, seed=9), dict(periodic=False))


def test_mixed_time_bins():
    ic = ics.uniform_random(24 ** 3, seed=21)
    rng = np.random.default_rng(1)
    ic["time_bin"] = rng.integers(1, 4, len(ic["h"])).astype(np.int8)
    _check(ic, dict(periodic=True, max_active_bin=2))
