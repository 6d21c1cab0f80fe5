#!/bin/bash
# N GPUs of one box: the NCCL slab tests, then bench.py under torch.distributed.run
cd "$(dirname "$0")/.."
N=${1:-2}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_sweep.py -x -q -m gpu -k developed_start > gpurun_out/t_dev.log 2>&1; echo "developed test rc=$?"; tail -5 gpurun_out/t_dev.log
python -m pytest tests/test_slabs.py -x -q -m gpu -k nccl > gpurun_out/t_nccl.log 2>&1; echo "nccl tests rc=$?"; tail -15 gpurun_out/t_nccl.log
for mode in 2 1; do
SLAB_SWEEP=$mode python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/dist_slab_check.py > gpurun_out/t_dist_mode$mode.log 2>&1; echo "dist check mode $mode rc=$?"; grep -c identical gpurun_out/t_dist_mode$mode.log; grep -i "differs\|error" gpurun_out/t_dist_mode$mode.log | head -5
done
for w in colliding_streams_adiabatic_varh_n64m colliding_streams_adiabatic_varh_n64m_pow2 colliding_streams_adiabatic_varh_developed_n64m; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 --workload $w > gpurun_out/bm_${N}_$w.json 2> gpurun_out/bm_${N}_$w.err; echo "bench $w x$N rc=$?"
done
timeout 600 python bench.py --steps 30 --warmup 5 --workload colliding_streams_adiabatic_varh_developed_n64m > gpurun_out/bm_1_developed.json 2> gpurun_out/bm_1_developed.err; echo "bench developed x1 rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bm_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f.split('/')[-1], 'value %.3e ms %.3f' % (d['value'], d['ms_per_step']), {k: round(v,3) for k,v in r['phase_ms'].items()}, r['kernel'], d['config'].get('sweep'), d['config'].get('parallelism'), 'parity', d.get('slab_parity'), 'e2e %.3e' % d['e2e']['value'])
    except Exception as e:
        print(f, 'ERR', e); print(open(f.replace('.json','.err')).read()[-1500:])
PY
