#!/bin/bash
# One GPU-box pass on a named workload: bench + (optionally) ncu --set full of the hot kernels.
# usage (through gpurun): bash tools/gpu_cfg.sh <tag> <workload> <reads> [noprof]
tag=$1; wl=$2; reads=$3
out=gpurun_out
mkdir -p $out
timeout 2400 python bench.py --workload $wl --reads $reads --steps 3 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 1500 $out/${tag}_bench.json; tail -5 $out/${tag}_bench.err
if [ "$4" != "noprof" ]; then
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_seed|k_cand|k_verify|k_trace|k_compact' \
  --launch-skip 14 -c 8 -o $out/${tag}_full -f python bench.py --workload $wl --reads $reads --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
