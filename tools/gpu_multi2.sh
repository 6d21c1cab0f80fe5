#!/bin/bash
cd "$(dirname "$0")/.."
N=${1:-2}; mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_slabs.py -x -q -m gpu -k nccl > gpurun_out/t_nccl.log 2>&1; echo "nccl tests rc=$?"; tail -3 gpurun_out/t_nccl.log
run() { tag=$1; shift; timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N "$@" > gpurun_out/m${N}_$tag.json 2> gpurun_out/m${N}_$tag.err; echo "$tag x$N rc=$?"; }
run strong --steps 30 --warmup 5
run weak64 --steps 20 --warmup 3 --scaling weak --e2e-steps 1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/m*_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d['roofline']
        print(f.split('/')[-1], 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), {k: round(v,3) for k,v in r['phase_ms'].items()}, d['config'].get('parallelism'), 'parity', d.get('slab_parity'), d['scaling'], d['config'].get('limit'), d['config'].get('n_particles'))
    except Exception as e:
        print(f, 'ERR', e); print(open(f.replace('.json','.err')).read()[-800:])
PY
