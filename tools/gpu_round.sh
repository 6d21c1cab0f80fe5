#!/bin/bash
# one GPU: whole GPU suite, default bench line, the labelled best case, the other BASELINE configs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "gpu tests rc=$?"
tail -12 gpurun_out/t_all.log
timeout 600 python bench.py > gpurun_out/b_default.json 2> gpurun_out/b_default.err; echo "bench default rc=$?"
for w in colliding_streams_adiabatic_varh_n64m_pow2 colliding_streams_adiabatic_varh_n16m colliding_streams_isothermal_fixedh_n1m colliding_streams_adiabatic_varh_n1m; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err; echo "bench $w rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/b_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f.split('/')[-1], 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), {k: round(v,3) for k,v in r['phase_ms'].items()}, r['kernel'], 'frac %.3f'%r['frac'], d['config'].get('sweep'), 'e2e %.3e' % d['e2e']['value'], 'clk', d['clocks'])
    except Exception as e:
        print(f, 'ERR', e)
PY
