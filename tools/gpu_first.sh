#!/bin/bash
# first GPU pass of a change: the sweep tests, then the whole GPU suite, then short bench lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_sweep.py -x -q -m gpu > gpurun_out/t_sweep.log 2>&1; echo "sweep tests rc=$?" 
tail -15 gpurun_out/t_sweep.log
python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_sweep.py > gpurun_out/t_all.log 2>&1; echo "all tests rc=$?"
tail -8 gpurun_out/t_all.log
for mode in 2 1; do
  timeout 300 python bench.py --workload colliding_streams_adiabatic_varh_n64000000 --steps 10 --warmup 3 --sweep-mode $mode --no-cpu-baseline > gpurun_out/b_gen64m_mode$mode.json 2> gpurun_out/b_gen64m_mode$mode.err; echo "bench mode $mode rc=$?"
done
timeout 300 python bench.py --workload colliding_streams_adiabatic_varh_n64m --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_64mi_auto.json 2> gpurun_out/b_64mi_auto.err; echo "bench 64Mi rc=$?"
timeout 300 python bench.py --workload colliding_streams_adiabatic_varh_n64m --steps 20 --warmup 5 --no-cpu-baseline --sweep-mode 2 > gpurun_out/b_64mi_serial.json 2> gpurun_out/b_64mi_serial.err; echo "bench 64Mi serial rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/b_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), d['roofline']['phase_ms'], d['config'].get('sweep'), 'e2e %.3e' % d['e2e']['value'])
    except Exception as e:
        print(f, 'ERR', e)
PY
