#!/bin/bash
# A/B of build-time variants of the kernels (sph_b200/variants/libsphb_<tag>.so) on one GPU
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for v in ${VARIANTS:-v1 v4 d5 d4}; do
  for w in colliding_streams_adiabatic_varh_n16m colliding_streams_adiabatic_varh_n64m colliding_streams_adiabatic_varh_n64m_pow2; do
    SPHB_LIB=$PWD/sph_b200/variants/libsphb_$v.so timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/var_${v}_$w.json 2> gpurun_out/var_${v}_$w.err
    TAG="$v $w" F=gpurun_out/var_${v}_$w.json python - <<'PY'
import json, os
d = json.loads(open(os.environ["F"]).read().strip().splitlines()[-1]); p = d["roofline"]["phase_ms"]
print("%s ms %.3f density %.3f force %.3f refine %.3f" % (os.environ["TAG"], d["ms_per_step"], p["density"], p["force_first"], p["force_refine"]))
PY
  done
done
