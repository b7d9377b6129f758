#!/bin/bash
# One GPU-box pass.  usage (through gpurun): bash tools/gpu_round.sh <tag> [stage ...]
#   stages (default: tests bench1m launches full):
#   tests     python -m pytest tests -m gpu            cand   only the candidate-stage parity tests
#   bench     the driver's command on the default workload (cfg3, 10 M reads), short step counts, no other workloads
#   bench1m   cfg3 on a 1 M-read file (what the ncu captures run on)
#   variants  bench1m under FEM_REG_VARIANT=0,1,2 and FEM_CAND_MODE=6 (k_cand_pair), resident leg only
#   ref       the reference arm on cfg3
#   launches  ncu launch list of the 1 M-read run       (after the same command exited 0 without ncu)
#   full      ncu --set full of the first full-size launch of each hot kernel, with source
#   probe     ncu --set full of the INT32 peak probe (k_int_peak)
#   others    bench.py with its other_workloads sub-runs (cfg2, cfg5, cfg4', cfg1)
#   e2etune   `FEM map` on the 10 M-read file under different host settings (copiers, readers, pre-allocation)
tag=${1:-rXX}; shift
stages=${*:-tests bench1m launches full}
out=gpurun_out
mkdir -p $out
{ nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv; nproc; uname -r; free -g | head -2; df -h /dev/shm /tmp | cat;
  cat /sys/kernel/mm/transparent_hugepage/shmem_enabled; } > $out/${tag}_box.txt 2>&1
small="--reads 1000000 --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 --others none"
summary() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("value %.1fM e2e %.1fM (%s)" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6, d["e2e"].get("form")), {k: round(v["ms"], 3) for k, v in d["kernels"].items()},
          "roofline %s %.3f seed %.3f" % (d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline_seed"]["frac"]),
          "int32 %.3f" % (d["roofline_verify_int32"].get("frac") or 0), "parity", (d.get("parity_vs_reference") or {}).get("identical"),
          "cpu", (d.get("cpu_baseline") or {}).get("value"), "MHz", d["clocks"].get("sm_mhz"))
    print("   e2e phases", (d["e2e"]["forms"].get("process_per_gpu") or {}).get("phases_s_rank0"), "single", (d["e2e"]["forms"].get("single_process") or {}).get("value"))
    for k, v in (d.get("other_workloads") or {}).items():
        print("  other", k, v.get("value"), (v.get("e2e") or {}).get("value"), (v.get("parity_vs_reference") or {}).get("identical"), v.get("error"))
except Exception as ex:
    print("unreadable", sys.argv[1], ex)
PY
}
for st in $stages; do case $st in
tests)
  timeout 1700 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest.log
  tail -3 $out/${tag}_pytest.log ;;
cand)
  timeout 1700 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k candidates > $out/${tag}_pytest_cand.log 2>&1; echo "pytest cand rc=$?" | tee -a $out/${tag}_pytest_cand.log
  tail -15 $out/${tag}_pytest_cand.log ;;
bench)
  timeout 1500 python bench.py --steps 5 --warmup 3 --others none > $out/${tag}_bench_cfg3.json 2> $out/${tag}_bench_cfg3.err; echo "bench rc=$?"
  summary $out/${tag}_bench_cfg3.json ;;
bench1m)
  timeout 900 python bench.py --reads 1000000 --steps 5 --warmup 3 --others none > $out/${tag}_bench_cfg3_1M.json 2> $out/${tag}_bench_cfg3_1M.err; echo "bench1m rc=$?"
  summary $out/${tag}_bench_cfg3_1M.json ;;
variants)
  for v in "FEM_REG_VARIANT=0" "FEM_REG_VARIANT=1" "FEM_REG_VARIANT=2" "FEM_CAND_MODE=6" $VARIANTS; do
    n=$(echo $v | tr '= ' '__')
    env $v timeout 600 python bench.py $small > $out/${tag}_var_$n.json 2> $out/${tag}_var_$n.err; echo "$v rc=$?"
    summary $out/${tag}_var_$n.json
  done ;;
others)
  timeout 1700 python bench.py --reads 1000000 --steps 3 --warmup 3 --no-cpu-baseline --others all > $out/${tag}_bench_others.json 2> $out/${tag}_bench_others.err; echo "others rc=$?"
  summary $out/${tag}_bench_others.json ;;
ref)
  timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench_ref.err; echo "ref rc=$?"
  head -c 300 $out/${tag}_bench_ref.json; echo ;;
launches)
  python bench.py $small > $out/${tag}_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_ncu_launches.csv \
    python bench.py $small > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?" ;;
full)
  # the first pass over the whole 1 M-read batch (femgpu_map_batch): one launch of each hot kernel
  python bench.py $small > $out/${tag}_plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_seed|k_cand_reg|k_verify|k_trace' -c 4 \
    -o $out/${tag}_full -f python bench.py $small > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?" ;;
probe)
  python bench.py $small > $out/${tag}_plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none -k regex:'k_int_peak' -c 6 \
    -o $out/${tag}_probe -f python bench.py $small > $out/${tag}_ncu_probe.log 2>&1; echo "ncu probe rc=$?" ;;
e2etune)
  bash tools/e2e_tune.sh $tag "FEM_SAM_COPIERS=10" "FEM_READERS=8 FEM_SAM_COPIERS=10" "FEM_SAM_PREALLOC=0" "FEM_SAM_MMAP=0" "FEM_SAM_POPULATE=0" $TUNES ;;
esac; done
ls -la $out | tail -20
