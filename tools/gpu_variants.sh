#!/bin/bash
# A/B of build-time variants of the kernels (sph_b200/variants/libsphb_v*.so) on one GPU
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in ${VARIANTS:-1 4}; do
  for w in colliding_streams_adiabatic_varh_n16m colliding_streams_adiabatic_varh_n64m colliding_streams_adiabatic_varh_n64m_pow2; do
    SPHB_LIB=$PWD/sph_b200/variants/libsphb_v$v.so timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/v${v}_$w.json 2> gpurun_out/v${v}_$w.err
    python -c "
import json;d=json.loads(open('gpurun_out/v${v}_$w.json').read().strip().splitlines()[-1]);p=d['roofline']['phase_ms'];print('v$v $w ms %.3f density %.3f force %.3f refine %.3f'%(d['ms_per_step'], p['density'], p['force_first'], p['force_refine']))"
  done
done
