#!/bin/bash
# compute-sanitizer over the hot path ("race detection"): smoke() plus one parity case per
# candidate path, seed-DP form and SAM path.  One tool per gpurun call (B200_PROFILING.md).
# usage (through gpurun): bash tools/gpu_sanitize.sh <tag> racecheck|memcheck|synccheck|initcheck
tag=${1:-rXX}; tool=${2:-racecheck}
out=gpurun_out
mkdir -p $out
# the same command must have passed without the tool first
sel='test_candidates_match_oracle or test_candidates_reg_path or test_candidates_big_path or test_seed_selection_forms_agree or test_device_sam_text_matches_oracle or test_mappings_match_oracle'
cases='cfg3_e7_s5 or rep_e7_s2 or dense_q9 or withN or cfg2_e5_s4'
python __graft_entry__.py smoke > $out/${tag}_san_plain.log 2>&1 &&
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "($sel) and ($cases)" >> $out/${tag}_san_plain.log 2>&1 || { echo "plain run failed"; tail -5 $out/${tag}_san_plain.log; exit 1; }
tail -2 $out/${tag}_san_plain.log
export FEM_SANITIZE=1
timeout 1700 compute-sanitizer --tool $tool --print-limit 20 --log-file $out/${tag}_${tool}_smoke.log python __graft_entry__.py smoke > $out/${tag}_${tool}_smoke.out 2>&1; echo "smoke under $tool rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" $out/${tag}_${tool}_smoke.log | tail -3
timeout 2400 compute-sanitizer --tool $tool --print-limit 20 --log-file $out/${tag}_${tool}_parity.log \
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "($sel) and ($cases)" > $out/${tag}_${tool}_parity.out 2>&1; echo "parity under $tool rc=$?"
tail -2 $out/${tag}_${tool}_parity.out
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" $out/${tag}_${tool}_parity.log | sort | uniq -c | tail -5
