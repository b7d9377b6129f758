#!/bin/bash
# device SAM text: tests, then the cfg2 bench with the text built on the device / on the host.
# usage: bash tools/gpu_sam.sh <tag> [pytest -k expression]
tag=$1; sel=${2:-}
out=gpurun_out; mkdir -p $out
if [ -n "$sel" ]; then timeout 1500 python -m pytest tests -m gpu -x -q -k "$sel" > $out/${tag}_pytest.log 2>&1
else timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; fi
echo "pytest rc=$?"; tail -25 $out/${tag}_pytest.log
for mode in 1 0; do
  FEM_DEVICE_SAM=$mode timeout 900 python bench.py --steps 5 --warmup 3 > $out/${tag}_cfg2_sam$mode.json 2> $out/${tag}_cfg2_sam$mode.err; echo "cfg2 sam=$mode rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("$out/${tag}_cfg2_sam$mode.json")); print("cfg2 sam=$mode", d["value"], d["e2e"], {k:round(v["ms"],3) for k,v in d["kernels"].items()}, d.get("parity_vs_reference",{}).get("identical"))
except Exception as ex: print("cfg2 parse failed", ex)
PY
  grep -i "phases\|e2e" $out/${tag}_cfg2_sam$mode.err | tail -4
done
