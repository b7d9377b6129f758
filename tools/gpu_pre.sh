#!/bin/bash
for v in "FEM_SAM_COPIERS=6" "FEM_SAM_COPIERS=8" "FEM_SAM_COPIERS=10" "FEM_SAM_COPIERS=13" "FEM_SAM_PREALLOC_PERCENT=60" "FEM_SAM_PREALLOC=0"; do
  echo "== $v"; env $v bash tools/gpu_trace.sh r02f | tail -4
done
