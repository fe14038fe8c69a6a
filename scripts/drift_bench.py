"""k_drift alone at side^3 (default 256): ms per launch and GB/s of its 284 algorithmic bytes per particle."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from swift_subtask_dev_b200 import HydroBackend, ics, make_params
side = int(sys.argv[1]) if len(sys.argv) > 1 else 256
ic = ics.uniform_random(side ** 3, seed=42)
ic["time_bin"] = np.zeros(side ** 3, dtype=np.int8)
os.environ["SWIFT_CUDA_PATH"] = "0"   # no neighbour lists: this script only drifts
be = HydroBackend(make_params())
be.build(ic)
be.integrator_init(cfl_condition=0.1, log_max_h_change=0.11, minimal_internal_energy=0.0, dt_min=1e-12, dt_max=1e-2,
                   time_base=1.0 / float(1 << 57))
ms = []
for k in range(1, 8):
    be.set_time(k << 40, 56)
    be.drift((k - 1) << 40)
    ms.append(be.integrator_timings()["drift_ms"])
be.close()
best = min(ms[2:])
print({"side": side, "drift_ms": ms, "best_ms": best, "gbs": 284 * side ** 3 / (best * 1e-3) / 1e9})
