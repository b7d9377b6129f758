#!/bin/bash
# Last pass of a round: all GPU tests, the bench, the ncu launch list, the reference arm (no ncu --set full).
# usage (through gpurun): bash tools/gpu_final.sh <tag>
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('$out/${tag}_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], {k: round(v['ms'], 3) for k, v in d['kernels'].items()}, d.get('parity_vs_reference',{}).get('identical'))"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_ncu_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench.err; echo "ref rc=$?"
