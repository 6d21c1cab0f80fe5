#!/bin/bash
# last one-GPU pass of the round: GPU suite (without the 16 M oracle test), two kernel variants, ncu launch list, sanitizer
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 240 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_parity.py::test_adiabatic_variable_h_at_sixteen_million > gpurun_out/t_all.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/t_all.log
VARIANTS="v4 d5" W="colliding_streams_adiabatic_varh_n16m colliding_streams_adiabatic_varh_n16m_pow2"
for v in cur $VARIANTS; do for w in $W; do
  L=$PWD/sph_b200/variants/libsphb_$v.so; [ $v = cur ] && L=$PWD/sph_b200/libsphb.so
  SPHB_LIB=$L timeout 100 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 0 > gpurun_out/var_${v}_$w.json 2> gpurun_out/var_${v}_$w.err
  TAG="$v $w" F=gpurun_out/var_${v}_$w.json python - <<'PY'
import json, os
try:
    d = json.loads(open(os.environ["F"]).read().strip().splitlines()[-1]); p = d["roofline"]["phase_ms"]
    print("%s ms %.3f density %.3f force %.3f refine %.3f" % (os.environ["TAG"], d["ms_per_step"], p["density"], p["force_first"], p["force_refine"]))
except Exception as e:
    print(os.environ["TAG"], "ERR", e)
PY
done; done
timeout 120 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list_plain.json 2> gpurun_out/ncu_list_plain.err && \
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_default.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 100 python tools/sanitize_case.py all > gpurun_out/sanitize_plain.log 2>&1; echo "sanitize plain rc=$?"; cat gpurun_out/sanitize_plain.log | tail -4
for tool in memcheck racecheck; do
  timeout 170 compute-sanitizer --tool $tool --print-limit 10 python tools/sanitize_case.py single > gpurun_out/sanitize_${tool}_single.log 2>&1; echo "$tool single rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|identical" gpurun_out/sanitize_${tool}_single.log | tail -3
done
