#!/bin/bash
# Last pass of a round: all GPU tests, the driver's bench command and the reference arm on the default workload.
# usage (through gpurun): bash tools/gpu_final.sh <tag>
tag=${1:-rXX}
bash tools/gpu_round.sh $tag tests
out=gpurun_out
timeout 1700 python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
head -c 1500 $out/${tag}_bench.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench.err; echo "ref rc=$?"
