#!/bin/bash
# all BASELINE configs + the north-star pair through the fused device path -> gpurun_out/r1c_workloads.jsonl
out=gpurun_out/r1c_workloads.jsonl
: > $out
run() { timeout 900 python tools/run_workload.py "$@" 2>gpurun_out/wl.err | tail -1 >> $out || echo "{\"failed\": \"$*\"}" >> $out; }
run --workload config1_tp53_5Mbp --check-popc --check-oracle-mbp 5
run --workload config2_tp53_chr17 --check-popc --check-oracle-mbp 20
run --workload config3_16S_500Mbp --check-popc --check-oracle-mbp 10
run --workload config3_16S_256var_500Mbp --check-popc --check-oracle-mbp 10
run --workload config5_sweep_1Gbp --primer-len 18 --k 0
run --workload config5_sweep_1Gbp --primer-len 18 --k 4 --check-oracle-mbp 5
run --workload config5_sweep_1Gbp --primer-len 25 --k 2 --check-popc
run --workload config5_sweep_1Gbp --primer-len 35 --k 4 --check-oracle-mbp 5
run --workload config5_sweep_1Gbp --primer-len 35 --k 2 --v16
run --workload northstar_256fold_3Gbp --check-oracle-mbp 5
run --workload config4_20plex_3Gbp --check-oracle-mbp 2
cat $out | cut -c1-420
