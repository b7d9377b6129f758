#!/bin/bash
# per-batch timeline of the end-to-end run (FEM_TRACE).  usage: bash tools/gpu_trace.sh <tag> [threads] [batch]
out=gpurun_out; mkdir -p $out
FEM_TRACE=1 timeout 600 python bench.py --steps 2 --warmup 3 --e2e-steps 3 --no-cpu-baseline --e2e-threads ${2:-0} --e2e-batch ${3:-0} > $out/$1_trace.json 2> $out/$1_trace.err
grep "trace" $out/$1_trace.err | tail -24
python - <<PY
import json
d=json.load(open("$out/$1_trace.json")); print(round(d["e2e"]["value"]), d["e2e"]["phases_s"])
PY
