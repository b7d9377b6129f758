#!/bin/bash
# One GPU-box pass for the k_cand_small rewrite: targeted parity first, A/B against the previous kernel
# (FEM_SMALL_SHAPE=9), then the whole round (tools/gpu_round.sh), then cfg5 (the wide shape).
# usage (through gpurun): bash tools/gpu_small_ab.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "small or candidates_default or arena or big_path" > $out/${tag}_pytest_small.log 2>&1; echo "pytest small rc=$?" | tee -a $out/${tag}_pytest_small.log
tail -3 $out/${tag}_pytest_small.log
FEM_SMALL_SHAPE=9 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_v1.json 2> $out/${tag}_bench_v1.err; echo "bench v1 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_new.json 2> $out/${tag}_bench_new.err; echo "bench new rc=$?"
python - <<PY
import json
for t in ("v1", "new"):
    try:
        d = json.loads(open("$out/${tag}_bench_%s.json" % t).read().strip().splitlines()[-1])
        print(t, "value %.1fM" % (d["value"] / 1e6), {k: round(v["ms"], 3) for k, v in d["kernels"].items()})
    except Exception as ex:
        print(t, "unreadable", ex)
PY
timeout 900 python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 400 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_ncu_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_seed|k_cand|k_verify|k_trace|k_compact|k_sam' \
  --launch-skip 24 -c 8 -o $out/${tag}_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 1200 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
timeout 900 python bench.py --workload cfg5 --reads 1000000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_cfg5.json 2> $out/${tag}_bench_cfg5.err; echo "cfg5 rc=$?"
tail -c 900 $out/${tag}_bench_cfg5.json
