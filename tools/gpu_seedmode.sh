#!/bin/bash
# k_seed forms on a workload.  usage: bash tools/gpu_seedmode.sh <workload> <reads>
out=gpurun_out; mkdir -p $out
for m in 1 3 0; do
  FEM_SEED_DP_MODE=$m timeout 1200 python bench.py --workload $1 --reads $2 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/seedmode_$m.json 2> $out/seedmode_$m.err
  python - <<PY
import json
d=json.load(open("$out/seedmode_$m.json")); print("$1 seed_dp_mode $m:", round(d["value"]), {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
PY
done
