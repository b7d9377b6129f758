#!/bin/bash
# Benches WITH the reference binary beside them (parity_vs_reference on the CPU sample): cfg1, cfg5, cfg2.
# usage (through gpurun): bash tools/gpu_parity_cfgs.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 600 python bench.py --workload cfg1 --steps 5 --warmup 3 > $out/${tag}_bench_cfg1.json 2> $out/${tag}_bench_cfg1.err; echo "cfg1 rc=$?"
timeout 1500 python bench.py --workload cfg5 --reads 1000000 --cpu-sample 100000 --steps 3 --warmup 3 > $out/${tag}_bench_cfg5.json 2> $out/${tag}_bench_cfg5.err; echo "cfg5 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 > $out/${tag}_bench_cfg2.json 2> $out/${tag}_bench_cfg2.err; echo "cfg2 rc=$?"
python - <<PY
import json
for t in ("cfg1", "cfg5", "cfg2"):
    try:
        d = json.loads(open("$out/${tag}_bench_%s.json" % t).read().strip().splitlines()[-1])
        print(t, "value %.1fM e2e %.1fM" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6), {k: round(v["ms"], 3) for k, v in d["kernels"].items()},
              d.get("parity_vs_reference"), d.get("cpu_baseline"), d["clocks"])
    except Exception as ex:
        print(t, "unreadable", ex)
PY
