#!/bin/bash
# e2e leg over host thread counts and batch sizes.  usage: bash tools/gpu_e2e2.sh <tag>
out=gpurun_out; mkdir -p $out
nproc; uname -r
for v in "0 0" "2 0" "4 0" "6 0" "8 0" "16 0" "0 65536" "0 262144" "6 65536"; do
  set -- $v
  timeout 600 python bench.py --steps 2 --warmup 3 --e2e-steps 6 --no-cpu-baseline --e2e-threads $1 --e2e-batch $2 > $out/e2e2.json 2>> $out/e2e2.err
  python - <<PY
import json
d=json.load(open("$out/e2e2.json")); print("threads $1 batch $2:", round(d["e2e"]["value"]), {k: round(v, 4) for k, v in d["e2e"]["phases_s"].items()}, {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
PY
done
