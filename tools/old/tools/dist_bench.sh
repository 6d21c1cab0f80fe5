#!/bin/bash
# tools/dist_bench.sh N [extra bench args]: the driver's multi-GPU launch line, summary only
N=$1; shift
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline "$@" 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N=%d %.4g pps  %.3f ms/step' % (d['n_gpus'], d['value'], d['ms_per_step']), {k: round(v,3) for k,v in d['roofline']['phase_ms'].items()}, d['config']['parallelism'], 'e2e=%.4g' % d['e2e']['value'])"
