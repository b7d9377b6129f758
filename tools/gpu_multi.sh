#!/bin/bash
# tests + cfg2 bench + cfg3 bench (no profiler).  usage: bash tools/gpu_multi.sh <tag> [cfg3 reads]
tag=$1; reads=${2:-2000000}
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $out/${tag}_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > $out/${tag}_cfg2_bench.json 2> $out/${tag}_cfg2_bench.err; echo "cfg2 rc=$?"
python - <<PY
import json
try:
    d=json.load(open("$out/${tag}_cfg2_bench.json")); print("cfg2", d["value"], d["e2e"]["value"], {k:round(v["ms"],3) for k,v in d["kernels"].items()}, d.get("parity_vs_reference",{}).get("identical"))
except Exception as ex: print("cfg2 parse failed", ex)
PY
tail -3 $out/${tag}_cfg2_bench.err
timeout 2400 python bench.py --workload cfg3 --reads $reads --steps 3 --warmup 3 > $out/${tag}_cfg3_bench.json 2> $out/${tag}_cfg3_bench.err; echo "cfg3 rc=$?"
python - <<PY
import json
try:
    d=json.load(open("$out/${tag}_cfg3_bench.json")); print("cfg3", d["value"], d["e2e"]["value"], {k:round(v["ms"],3) for k,v in d["kernels"].items()}, d.get("parity_vs_reference",{}).get("identical"), d["roofline"]["frac"])
except Exception as ex: print("cfg3 parse failed", ex)
PY
tail -4 $out/${tag}_cfg3_bench.err
