#!/bin/bash
# cfg3 (3.1 Gbp, 150 bp, e=7): bench on <reads> reads and one ncu --set full capture of the candidate kernel with source.
# usage (through gpurun): bash tools/gpu_cfg3_ncu.sh <tag> <reads>
tag=$1; reads=${2:-1000000}
out=gpurun_out
mkdir -p $out
timeout 900 python bench.py --workload cfg3 --reads $reads --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_cfg3.json 2> $out/${tag}_bench_cfg3.err; echo "bench rc=$?"
tail -c 1200 $out/${tag}_bench_cfg3.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_cand' \
  --launch-skip 4 -c 1 -o $out/${tag}_cfg3_cand -f python bench.py --workload cfg3 --reads $reads --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_cfg3.log 2>&1; echo "ncu rc=$?"
timeout 900 python bench.py --workload cfg5 --reads 1000000 --steps 3 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_bench_cfg5.json 2> $out/${tag}_bench_cfg5.err; echo "cfg5 rc=$?"
tail -c 1500 $out/${tag}_bench_cfg5.json | head -c 700
