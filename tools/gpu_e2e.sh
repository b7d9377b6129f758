#!/bin/bash
# e2e leg only, variants by environment.  usage: bash tools/gpu_e2e.sh <tag>
out=gpurun_out; mkdir -p $out
for v in "FEM_SAM_POPULATE=1" "FEM_SAM_POPULATE=0" "FEM_SAM_MMAP=0"; do
  env $v timeout 600 python bench.py --steps 2 --warmup 3 --e2e-steps 6 --no-cpu-baseline > $out/$1_e2e.json 2>> $out/$1_e2e.err
  python - <<PY
import json
d=json.load(open("$out/$1_e2e.json")); print("$v", round(d["e2e"]["value"]), d["e2e"]["phases_s"])
PY
done
