#!/bin/bash
# Quick GPU-box pass: parity tests of the kernels, then cfg2 / cfg3 / cfg5 benches without the CPU baseline.
# usage (through gpurun): bash tools/gpu_quick3.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -x -q 2>&1 | tail -3
for wl in cfg2 cfg3 cfg5; do
  timeout 900 python bench.py --workload $wl --reads 1000000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "$wl rc=$?"
done
python - <<PY
import json
for t in ("cfg2", "cfg3", "cfg5"):
    try:
        d = json.loads(open("$out/${tag}_bench_%s.json" % t).read().strip().splitlines()[-1])
        print(t, "value %.1fM" % (d["value"] / 1e6), {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
    except Exception as ex:
        print(t, "unreadable", ex)
PY
