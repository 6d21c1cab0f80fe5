#!/bin/bash
# compute-sanitizer memcheck and racecheck over the small-N cases of tools/sanitize_case.py -> gpurun_out/sanitize_*.log
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/sanitize_case.py all > gpurun_out/sanitize_plain.log 2>&1; echo "plain rc=$?"
for tool in memcheck racecheck; do
  for c in single slabs; do
    timeout 1200 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py $c > gpurun_out/sanitize_${tool}_$c.log 2>&1
    echo "$tool $c rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|identical|n_ghost" gpurun_out/sanitize_${tool}_$c.log
  done
done
