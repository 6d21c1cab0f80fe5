#!/bin/bash
# scaling lines at N GPUs: strong (64 M in total, lattice + developed), weak (256 M per GPU, lattice)
cd "$(dirname "$0")/.."
N=${1:-8}
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N "$@" > gpurun_out/sc_${N}_$tag.json 2> gpurun_out/sc_${N}_$tag.err; echo "$tag x$N rc=$?"; }
run strong_lattice --steps 50 --warmup 5
run strong_developed --steps 30 --warmup 5 --workload colliding_streams_adiabatic_varh_developed_n64m
run strong_pow2 --steps 30 --warmup 5 --workload colliding_streams_adiabatic_varh_n64m_pow2
if [ "${2:-}" = "weak" ]; then
run weak_lattice_256m --steps 10 --warmup 3 --scaling weak --workload colliding_streams_adiabatic_varh_n256m
run weak_developed_256m --steps 10 --warmup 3 --scaling weak --workload colliding_streams_adiabatic_varh_developed_n256m
fi
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/sc_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f.split('/')[-1], 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), {k: round(v,3) for k,v in r['phase_ms'].items()}, d['config'].get('sweep'), d['config'].get('parallelism'), 'parity', d.get('slab_parity'), 'e2e %.3e' % d['e2e']['value'], d['scaling'])
    except Exception as e:
        print(f, 'ERR', e); print(open(f.replace('.json','.err')).read()[-1500:])
PY
