#!/bin/bash
# k_seed (one read-strand per warp: cfg3, cfg5): parity tests, cfg3 / cfg5 benches, one ncu --set full capture of k_seed on cfg5.
# usage (through gpurun): bash tools/gpu_seed_ncu.sh <tag> [ncu]
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py -x -q 2>&1 | tail -3
for wl in cfg3 cfg5; do
  timeout 900 python bench.py --workload $wl --reads 1000000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_$wl.json 2> $out/${tag}_bench_$wl.err; echo "$wl rc=$?"
done
python - <<PY
import json
for t in ("cfg3", "cfg5"):
    try:
        d = json.loads(open("$out/${tag}_bench_%s.json" % t).read().strip().splitlines()[-1])
        print(t, "value %.1fM" % (d["value"] / 1e6), {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
    except Exception as ex:
        print(t, "unreadable", ex)
PY
if [ "$2" == "ncu" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_seed' \
  --launch-skip 4 -c 1 -o $out/${tag}_cfg5_seed -f python bench.py --workload cfg5 --reads 1000000 --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_cfg5.log 2>&1; echo "ncu rc=$?"
fi
