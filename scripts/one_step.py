"""One step of a uniform box through the kernels SWIFT_CUDA_PATH selects (target of ncu captures).
  python scripts/one_step.py SIDE [steps]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from swift_subtask_dev_b200 import HydroBackend, make_params, ics
side = int(sys.argv[1]); steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
be = HydroBackend(make_params(periodic=True))
be.build(ics.uniform_random(side ** 3, seed=42))
for k in range(steps):
    be.step(do_sort=True, restore=True)
    print(be.timings(), be.counters())
be.close()
