#!/bin/bash
# ncu --set full of kernels matching a regex in the cfg2 bench.  usage: bash tools/gpu_ncu_k.sh <tag> <regex> [skip] [count] [workload] [reads]
tag=$1; re=$2; skip=${3:-0}; cnt=${4:-2}; wl=${5:-cfg2}; reads=${6:-0}
out=gpurun_out; mkdir -p $out
timeout 900 python bench.py --workload $wl --reads $reads --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 2 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json")); print(d["value"], d["e2e"]["value"], {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
PY
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$re" --launch-skip $skip -c $cnt -o $out/${tag}_full -f \
  python bench.py --workload $wl --reads $reads --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
