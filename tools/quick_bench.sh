#!/bin/bash
# quick timing line: value, ms/step, per-phase ms, step fraction (B_ns accounting)
python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.4g pps  %.3f ms/step' % (d['value'], d['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['phase_ms'].items()}, 'frac=%.3f' % d['roofline']['frac'], 'e2e=%.4g' % d['e2e']['value'])"
