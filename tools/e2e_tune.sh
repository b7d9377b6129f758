#!/bin/bash
# `FEM map` on the bench's 10 M-read cfg3 file under different host settings; prints the Time: and [phases] lines.
# usage (through gpurun): bash tools/e2e_tune.sh <tag> ["VAR=v ..." ...]
tag=${1:-rXX}; shift
out=gpurun_out; mkdir -p $out
d=/tmp/fem_bench/ref3100000000_rec24_rep0_seed3
if [ ! -f $d/reads_10000000_l150_e7.fq ]; then
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --others none --no-single-process > $out/${tag}_tune_prep.json 2> $out/${tag}_tune_prep.err
  echo "prep rc=$?"; ls -la $d | head
fi
fq=$d/reads_10000000_l150_e7.fq
for v in "X=0" "$@"; do
  for rep in 1 2; do
    rm -f /dev/shm/tune.sam
    env $v fem_b200/bin/FEM map -e 7 -t 13 --ref $d/ref.fa --index $d/index_k12_s5 --read1 $fq -o /dev/shm/tune.sam > $out/${tag}_tune.log 2>&1
    echo "rc=$? [$v] $(grep -E '^Time' $out/${tag}_tune.log) $(grep -E 'phases' $out/${tag}_tune.log | cut -c1-150)"
  done
done 2>&1 | tee $out/${tag}_e2etune.txt
ls -la /dev/shm/tune.sam; rm -f /dev/shm/tune.sam
