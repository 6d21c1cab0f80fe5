#!/bin/bash
# whole GPU suite with timings of the slowest tests
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu --durations=12 > gpurun_out/t_all.log 2>&1; echo "gpu tests rc=$?"
tail -40 gpurun_out/t_all.log
for w in colliding_streams_isothermal_fixedh_n1m colliding_streams_adiabatic_varh_n1m; do
  timeout 300 python bench.py --workload $w --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err; echo "bench $w rc=$?"
  python -c "
import json;d=json.loads(open('gpurun_out/b_$w.json').read().strip().splitlines()[-1]);print('$w', d['ms_per_step'], d['roofline']['phase_ms'], d['config']['sweep'])"
done
