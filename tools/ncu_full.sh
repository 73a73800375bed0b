#!/bin/bash
# ONE `ncu --set full` capture per call (plain run of the same command first).
#   bash tools/ncu_full.sh <tag> <skip> <count>   -> gpurun_out/<tag>.ncu-rep
tag=${1:-r2i_c4}; skip=${2:-12}; count=${3:-3}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --in-flight 1 --host-threads 1"
K='regex:dp_jit_scan|dp_duplex_kernel|dp_dimer_gather_kernel|dp_scan_bitslice_kernel'
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s $skip -c $count -o gpurun_out/${tag} -f $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/${tag}_ncu.log; ls -la gpurun_out/${tag}.ncu-rep
