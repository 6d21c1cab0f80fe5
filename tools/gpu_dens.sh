#!/bin/bash
# density kernel at 6 / 5 / 4 CTAs per SM (40 / 48 / 64 registers): same-box A/B
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in colliding_streams_adiabatic_varh_n16m colliding_streams_adiabatic_varh_n64m; do for v in cur d5 d4; do
  L=$PWD/sph_b200/variants/libsphb_$v.so; [ $v = cur ] && L=$PWD/sph_b200/libsphb.so
  SPHB_LIB=$L timeout 60 python bench.py --workload $w --steps 15 --warmup 3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/dens_${v}_$w.json 2> gpurun_out/dens_${v}_$w.err
  TAG="$v $w" F=gpurun_out/dens_${v}_$w.json python - <<'PY'
import json, os
try:
    d = json.loads(open(os.environ["F"]).read().strip().splitlines()[-1]); p = d["roofline"]["phase_ms"]
    print("%s ms %.3f density %.3f force %.3f" % (os.environ["TAG"], d["ms_per_step"], p["density"], p["force_first"]))
except Exception as e:
    print(os.environ["TAG"], "ERR", e)
PY
done; done
