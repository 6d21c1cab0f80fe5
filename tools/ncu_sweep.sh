#!/bin/bash
# ncu --set full capture of the serial sweep kernel (and the density kernel) on the generic 16M lattice
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
W=${1:-colliding_streams_adiabatic_varh_n16000000}
python bench.py --workload $W --steps 4 --warmup 3 --sweep-mode 2 --no-cpu-baseline > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_force_sweep|k_density' --launch-skip 8 --launch-count 4 \
    -o gpurun_out/sweep_full -f python bench.py --workload $W --steps 4 --warmup 3 --sweep-mode 2 --no-cpu-baseline > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"
ls -la gpurun_out/
