"""Wall-clock time of the host builder (tree, tasks, per-leaf interaction lists, lanes: csrc/host_tree.cpp) on this
machine's cores, stage by stage, without a GPU:  python scripts/host_plan_time.py [side] (SWIFT_CUDA_HOST_THREADS=1
for the single-threaded form).  The digest at the end is what tests/test_cabi.py compares between the two."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["SWIFT_CUDA_HOST_TIMING"] = "1"

import parity  # noqa: E402
from swift_subtask_dev_b200 import backend as B  # noqa: E402
from swift_subtask_dev_b200 import ics  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 128
ic = ics.uniform_random(side ** 3, seed=42)
prm = parity.gpu_params(dict(periodic=True))
t0 = time.time()
d = B.host_plan_digest(prm, ic)
print("side %d, %d particles, host threads %s: %.3f s including the digest; digest %016x" % (
    side, side ** 3, os.environ.get("SWIFT_CUDA_HOST_THREADS", "all"), time.time() - t0, d))
