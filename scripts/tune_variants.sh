#!/bin/bash
# Rebuilds the library on the GPU box with different launch bounds and prints the phase times of a short bench.
# usage: scripts/tune_variants.sh "D F" "D F" ...
out=gpurun_out/tune.log; : > $out
for v in "$@"; do set -- $v
  SWIFT_CUDA_NVCC_EXTRA="-DSWIFT_MINB_DENSITY=$1 -DSWIFT_MINB_FORCE=$2" python -c "
from swift_subtask_dev_b200 import build_library
build_library(force=True)" || exit 1
  python bench.py --steps 2 --warmup 1 --no-cpu 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('D$1 F$2', round(j['ms_per_step'],1), j['detail']['phase_ms'])" | tee -a $out
done
