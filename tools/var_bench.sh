#!/bin/bash
# bench.py's resident leg on 1 M reads under environment variants.  usage (through gpurun):
#   [READS=n] [E2E_STEPS=n] bash tools/var_bench.sh <tag> <workload> "<VAR=v ...>" ["<VAR=v ...>" ...]
tag=$1; wl=$2; shift 2
out=gpurun_out; mkdir -p $out
small="--reads ${READS:-1000000} --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps ${E2E_STEPS:-1} --others none --no-single-process"
for v in "$@"; do
  n=$(echo $v | tr '= ' '__')
  env $v timeout 600 python bench.py --workload $wl $small > $out/${tag}_${wl}_$n.json 2> $out/${tag}_${wl}_$n.err; echo "$wl [$v] rc=$?"
  python - $out/${tag}_${wl}_$n.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  value %.2fM e2e %.2fM" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6), {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
    print("  e2e phases", d["e2e"]["forms"]["process_per_gpu"].get("phases_s_rank0"))
except Exception as ex:
    print("  unreadable", ex)
PY
  grep 'copiers:' $out/${tag}_${wl}_$n.err | tail -3
done
