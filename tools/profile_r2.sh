#!/bin/bash
# Round-2 profiling call (one gpurun call): plain run of the exact command first, then the launch list and
# `ncu --set full` of the kernels of one step of config 4 and of the north-star pair.
#   bash tools/profile_r2.sh <tag>      -> gpurun_out/<tag>_launches.csv, <tag>_c4.ncu-rep, <tag>_ns.ncu-rep
tag=${1:-r2b}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --in-flight 1 --host-threads 1"
K='regex:dp_jit_scan|dp_duplex_kernel|dp_dimer_gather_kernel|dp_scan_bitslice_kernel'
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s 12 -c 3 -o gpurun_out/${tag}_c4 -f $CMD > gpurun_out/${tag}_ncu2.log 2>&1
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s 36 -c 3 -o gpurun_out/${tag}_ns -f $CMD > gpurun_out/${tag}_ncu3.log 2>&1
ls -la gpurun_out/ | tail -12
tail -3 gpurun_out/${tag}_ncu2.log gpurun_out/${tag}_ncu3.log
