#!/bin/bash
# the end-to-end leg on four cores (what a rank gets with 8 GPUs on a 32-core host).  usage: bash tools/gpu_small_host.sh
out=gpurun_out; mkdir -p $out
for v in "1 0 1" "2 0 2" "2 1 2" "2 0 1" "3 0 2" "3 1 2" "2 0 3"; do
  set -- $v
  export FEM_BLOCKING_SYNC=$2 FEM_READERS=$3
  taskset -c 0-3 timeout 600 python bench.py --steps 2 --warmup 3 --e2e-steps 4 --no-cpu-baseline --e2e-threads $1 > $out/small.json 2> $out/small.err
  python - <<PY
import json
d=json.load(open("$out/small.json")); print("4 cores, -t $1, blocking $2, readers $3:", round(d["e2e"]["value"]), {k: round(v, 3) for k, v in d["e2e"]["phases_s"].items()})
PY
done
