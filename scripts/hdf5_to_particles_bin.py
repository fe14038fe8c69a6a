"""Converts the gas particles of a SWIFT HDF5 initial-conditions file or snapshot (PartType0: Coordinates,
Velocities, Masses, SmoothingLengths, InternalEnergies, ParticleIDs, optional MaterialIDs, Densities, Pressures;
Header/BoxSize, Header/Time -- the field set of src/hydro/Planetary/hydro_io.h:55-84, 190-228) into the flat binary
particle file of include/swift_cuda.h ("Flat binary particle files"):

    python scripts/hdf5_to_particles_bin.py in.hdf5 out.bin

Needs h5py, which this repository's image does not have: the script has NOT been run here (the binary side it calls,
backend.write_particles, is covered by tests/test_particle_io.py).  Values are copied as they are: the file's
internal units are the run's."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

FIELDS = {"x": "Coordinates", "v": "Velocities", "mass": "Masses", "h": "SmoothingLengths", "u": "InternalEnergies",
          "id": "ParticleIDs", "mat_id": "MaterialIDs", "rho": "Densities", "a": "Accelerations", "pressure": "Pressures"}


def main():
    if len(sys.argv) != 3:
        print(__doc__)
        return 2
    try:
        import h5py
    except ImportError:
        print("h5py is not installed", file=sys.stderr)
        return 1
    from swift_subtask_dev_b200 import backend as B
    with h5py.File(sys.argv[1], "r") as f:
        gas = f["PartType0"]
        ic = {key: (np.asarray(gas[name]) if name in gas else None) for key, name in FIELDS.items()}
        box = np.atleast_1d(np.asarray(f["Header"].attrs["BoxSize"], dtype=np.float64))
        dim = tuple(box) if len(box) == 3 else (float(box[0]),) * 3
        time = float(np.atleast_1d(f["Header"].attrs.get("Time", 0.0))[0])
    B.write_particles(sys.argv[2], ic, dim, time=time)
    print("wrote %d particles to %s" % (len(ic["h"]), sys.argv[2]))
    return 0


if __name__ == "__main__":
    sys.exit(main())
