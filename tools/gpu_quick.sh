#!/bin/bash
# Quick GPU-box pass: candidate parity tests, cfg2 bench (no CPU baseline), optional extra bench commands.
# usage (through gpurun): bash tools/gpu_quick.sh <tag> [ncu]
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "small or candidates_default or arena or big_path" > $out/${tag}_pytest_small.log 2>&1; echo "pytest small rc=$?" | tee -a $out/${tag}_pytest_small.log
tail -3 $out/${tag}_pytest_small.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_new.json 2> $out/${tag}_bench_new.err; echo "bench new rc=$?"
FEM_CAND_MODE=4 timeout 900 python bench.py --workload cfg5 --reads 1000000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_cfg5_small.json 2> $out/${tag}_bench_cfg5_small.err; echo "cfg5 small rc=$?"
python - <<PY
import json
for t in ("new", "cfg5_small"):
    try:
        d = json.loads(open("$out/${tag}_bench_%s.json" % t).read().strip().splitlines()[-1])
        print(t, "value %.1fM" % (d["value"] / 1e6), {k: round(v["ms"], 3) for k, v in d["kernels"].items()}, d["stats"])
    except Exception as ex:
        print(t, "unreadable", ex)
PY
if [ "$2" == "ncu" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_seed|k_cand|k_trace' \
  --launch-skip 24 -c 8 -o $out/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
