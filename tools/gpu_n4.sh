#!/bin/bash
# N-GPU bench under torchrun (ours + reference arm).  usage: bash tools/gpu_n4.sh <tag> [N]
tag=$1; N=${2:-4}
out=gpurun_out; mkdir -p $out
nproc; nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > $out/${tag}_bench_n$N.json 2> $out/${tag}_bench_n$N.err; echo "bench N=$N rc=$?"
python - <<PY
import json
d=json.load(open("$out/${tag}_bench_n$N.json")); print(d["n_gpus"], d["value"], d["e2e"]["value"], d["ms_per_step"], d["clocks"], d["e2e"].get("host_threads"))
PY
timeout 600 $TR bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $out/${tag}_ref_n$N.json 2>> $out/${tag}_bench_n$N.err; echo "ref N=$N rc=$?"
cut -c1-160 $out/${tag}_ref_n$N.json
