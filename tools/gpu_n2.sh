#!/bin/bash
# 2-GPU pass: bench under torchrun (ours + reference arm), the one-process-per-GPU `FEM map` driver vs the
# single-process binary, and FEM map --gpus 2.  usage: bash tools/gpu_n2.sh <tag> [N]
tag=$1; N=${2:-2}
out=gpurun_out; mkdir -p $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "bench N=$N rc=$?"
tail -c 700 $out/${tag}_bench_n$N.json; echo
timeout 600 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $out/${tag}_ref_n$N.json 2>> $out/${tag}_bench_n$N.err; echo "ref N=$N rc=$?"
cut -c1-200 $out/${tag}_ref_n$N.json
# sharded driver vs single process
d=/tmp/fem_n2; mkdir -p $d
tools/femsim ref 7 20000000 5 20 $d/ref.fa && tools/femsim reads 1007 $d/ref.fa 300001 100 4 $d/reads.fq
fem_b200/bin/FEM index 12 4 $d/ref.fa $d/idx --gpu 2> /dev/null
fem_b200/bin/FEM map -e 4 -t 8 --ref $d/ref.fa --index $d/idx --read1 $d/reads.fq -o $d/one.sam 2> $d/one.err; echo "single rc=$?"
fem_b200/bin/FEM map -e 4 -t 8 --gpus $N --ref $d/ref.fa --index $d/idx --read1 $d/reads.fq -o $d/multi.sam 2> $d/multi.err; echo "--gpus $N rc=$?"
timeout 600 $TR -m fem_b200.shard -e 4 -t 4 --ref $d/ref.fa --index $d/idx --read1 $d/reads.fq -o $d/shard.sam 2> $d/shard.err; echo "shard rc=$?"
FEM_CLONE_RESIDENT=0 fem_b200/bin/FEM map -e 4 -t 8 --gpus $N --ref $d/ref.fa --index $d/idx --read1 $d/reads.fq -o $d/multi0.sam 2> $d/multi0.err; echo "--gpus $N, every device loads for itself rc=$?"
cmp $d/one.sam $d/multi0.sam && echo "FEM map --gpus $N (no clone) == single"
grep -i "resident\|load" $d/multi.err $d/multi0.err | head
cmp $d/one.sam $d/multi.sam && echo "FEM map --gpus $N == single"; cmp $d/one.sam $d/shard.sam && echo "torchrun shard driver == single"
grep "number of" $d/one.err | tr '\n' ';'; echo; grep "number of\|Time" $d/shard.err | tr '\n' ';'; echo
