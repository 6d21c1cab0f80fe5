#!/bin/bash
# last one-GPU pass of the round: GPU suite (without the 16 M oracle test), ncu launch list of the default bench, sanitizer
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 240 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_parity.py::test_adiabatic_variable_h_at_sixteen_million > gpurun_out/t_all.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/t_all.log
timeout 100 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list_plain.json 2> gpurun_out/ncu_list_plain.err && \
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_default.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
python -c "
import json;d=json.loads(open('gpurun_out/ncu_list_plain.json').read().strip().splitlines()[-1]);print('default ms %.3f'%d['ms_per_step'], d['roofline']['phase_ms'], 'frac %.3f'%d['roofline']['frac'])"
for tool in memcheck racecheck; do
  timeout 120 compute-sanitizer --tool $tool --print-limit 10 python tools/sanitize_case.py single > gpurun_out/sanitize_${tool}_single.log 2>&1; echo "$tool single rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|identical" gpurun_out/sanitize_${tool}_single.log | tail -3
done
