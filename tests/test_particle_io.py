"""Flat binary particle files (the second half of SURVEY.md 8(f) N4; include/swift_cuda.h): the field set of the
reference's HDF5 initial conditions and snapshots for gas (src/hydro/Planetary/hydro_io.h:55-84, 190-228) as one
file.  Host only: no GPU, no HDF5.  There is no reference output to compare with (HDF5 is absent here, and the
format is this library's own): round trips, the layout on disk, and what is refused."""
import os
import struct

import numpy as np
import pytest

from swift_subtask_dev_b200 import backend as B
from swift_subtask_dev_b200 import ics


def _ic(n=3000):
    ic = ics.two_body_impact(n, seed=3)
    dim = ic.pop("dim")
    ic.pop("h_max")
    return ic, dim


def test_round_trip_initial_conditions(tmp_path):
    ic, dim = _ic()
    path = str(tmp_path / "ic.bin")
    B.write_particles(path, ic, dim, time=0.25)
    back, dim2, time = B.read_particles(path)
    assert dim2 == tuple(float(v) for v in dim) and time == 0.25
    for key in ("x", "v", "mass", "h", "u", "id", "mat_id"):
        assert back[key].dtype == np.asarray(ic[key]).dtype and np.array_equal(back[key], ic[key]), key
    assert back["rho"] is None and back["a"] is None and back["pressure"] is None
    # on disk: 128-byte header, then the arrays in hydro_io.h's order, each padded to 16 bytes
    n = len(ic["h"])
    pad = lambda b: (b + 15) // 16 * 16
    assert os.path.getsize(path) == 128 + pad(24 * n) + pad(12 * n) + 3 * pad(4 * n) + pad(8 * n) + pad(4 * n)
    blob = open(path, "rb").read()
    assert blob[:8] == b"SWCUPRT1"
    assert struct.unpack_from("<q3dd2i", blob, 8) == (n, *[float(v) for v in dim], 0.25, 63 | 64, 1)
    assert np.array_equal(np.frombuffer(blob, dtype="<f8", count=3 * n, offset=128).reshape(n, 3), ic["x"])


def test_round_trip_snapshot_fields_and_other_schemes(tmp_path):
    ic = ics.uniform_random(1001, seed=4)            # no MaterialIDs: Minimal / SPHENIX
    rng = np.random.default_rng(0)
    ic["rho"] = rng.random(1001).astype(np.float32)
    ic["a"] = rng.standard_normal((1001, 3)).astype(np.float32)
    ic["pressure"] = rng.random(1001).astype(np.float32)
    path = str(tmp_path / "snap.bin")
    B.write_particles(path, ic, (1.0, 1.0, 1.0))
    back, dim, time = B.read_particles(path)
    assert back["mat_id"] is None
    for key in ("x", "v", "mass", "h", "u", "id", "rho", "a", "pressure"):
        assert np.array_equal(back[key], ic[key]), key


def test_empty_file_and_refusals(tmp_path):
    ic, dim = _ic(300)
    empty = {k: (np.asarray(v)[:0] if v is not None else None) for k, v in ic.items()}
    path = str(tmp_path / "empty.bin")
    B.write_particles(path, empty, dim)
    back, _, _ = B.read_particles(path)
    assert len(back["h"]) == 0 and os.path.getsize(path) == 128
    # a compulsory field is missing
    part = dict(ic)
    part["u"] = None
    with pytest.raises(B.SwiftCudaError, match="compulsory"):
        B.write_particles(str(tmp_path / "x.bin"), part, dim)
    # truncated, extended, foreign
    good = str(tmp_path / "good.bin")
    B.write_particles(good, ic, dim)
    blob = open(good, "rb").read()
    for name, data in (("short.bin", blob[:-16]), ("long.bin", blob + b"\0" * 16), ("foreign.bin", b"HDF5" + blob[4:])):
        p = str(tmp_path / name)
        open(p, "wb").write(data)
        with pytest.raises(B.SwiftCudaError):
            B.read_particles(p)
    with pytest.raises(B.SwiftCudaError, match="cannot open"):
        B.read_particles(str(tmp_path / "nope.bin"))
