#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 1500 python bench.py --workload cfg5 --reads 1000000 --steps 3 --warmup 3 --cpu-sample 100000 > $out/$1_cfg5_bench.json 2> $out/$1_cfg5_bench.err; echo "cfg5 rc=$?"; tail -3 $out/$1_cfg5_bench.err
rm -rf /tmp/fem_bench
timeout 2400 python bench.py --workload cfg4p --reads 500000 --steps 3 --warmup 3 --cpu-sample 50000 > $out/$1_cfg4p_bench.json 2> $out/$1_cfg4p_bench.err; echo "cfg4p rc=$?"; tail -3 $out/$1_cfg4p_bench.err
python - <<PY
import json
for w in ("cfg5","cfg4p"):
    try:
        d=json.load(open("$out/$1_%s_bench.json"%w)); print(w, d["value"], d["e2e"]["value"], {k:round(v["ms"],3) for k,v in d["kernels"].items()}, d.get("parity_vs_reference"), d.get("cpu_baseline",{}).get("value"), d["stats"], d["gpu_launches"])
    except Exception as ex: print(w, "parse failed", ex)
PY
