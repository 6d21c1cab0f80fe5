#!/bin/bash
# build-time variants of the serial sweep kernel (sph_b200/variants/libsphb_v*.so), 16M and 64M generic lattice; 256M single GPU
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in 0 1 2 3; do
  for w in colliding_streams_adiabatic_varh_n16m colliding_streams_adiabatic_varh_n64m; do
    SPHB_LIB=$PWD/sph_b200/variants/libsphb_v$v.so timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --sweep-mode 2 --no-cpu-baseline > gpurun_out/v${v}_$w.json 2> gpurun_out/v${v}_$w.err
    python -c "
import json;d=json.loads(open('gpurun_out/v${v}_$w.json').read().strip().splitlines()[-1]);print('v$v $w ms %.3f force %.3f refine %.3f'%(d['ms_per_step'], d['roofline']['phase_ms']['force_first'], d['roofline']['phase_ms']['force_refine']))"
  done
done
timeout 900 python bench.py --workload colliding_streams_adiabatic_varh_n256m --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_256m.json 2> gpurun_out/b_256m.err; echo "256M rc=$?"
python -c "
import json;d=json.loads(open('gpurun_out/b_256m.json').read().strip().splitlines()[-1]);print('256M ms %.3f'%d['ms_per_step'], d['roofline']['phase_ms'], d['config']['sweep'], 'e2e %.3e'%d['e2e']['value'])"
tail -3 gpurun_out/b_256m.err
