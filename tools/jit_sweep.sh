#!/bin/bash
# Block size / occupancy / plane-mode variants of the run-time specialised scan kernel on the two 3.1 Gbp
# workloads -> gpurun_out/jit_sweep.jsonl  (ms_main_kernel is the CUDA-event time of the scan kernel)
out=gpurun_out/jit_sweep.jsonl
: > $out
run() {
  local tag="$1"; shift
  local line
  line=$(timeout 600 python tools/run_workload.py "$@" 2>gpurun_out/sweep.err | tail -1)
  if [ -z "$line" ]; then line="{\"failed\": \"$*\"}"; tail -3 gpurun_out/sweep.err; fi
  echo "{\"tag\": \"$tag\", \"r\": $line}" >> $out
}
for cfg in "256 1" "256 2" "128 2" "128 3" "128 4"; do
  set -- $cfg
  DEGENPRIMER_B200_JIT_THREADS=$1 DEGENPRIMER_B200_JIT_MIN_BLOCKS=$2 run "c4 shared t$1 b$2" --workload config4_20plex_3Gbp --algo 3 --check-oracle-mbp 2
done
DEGENPRIMER_B200_JIT_SHARED_FROM=100 DEGENPRIMER_B200_JIT_MIN_BLOCKS=3 run "c4 folded t256 b3" --workload config4_20plex_3Gbp --algo 3 --check-oracle-mbp 2
run "c4 generic" --workload config4_20plex_3Gbp --algo 2 --no-jit
for cfg in "256 3" "256 2" "128 4" "128 6"; do
  set -- $cfg
  DEGENPRIMER_B200_JIT_THREADS=$1 DEGENPRIMER_B200_JIT_MIN_BLOCKS=$2 run "ns folded t$1 b$2" --workload northstar_256fold_3Gbp --algo 3 --check-oracle-mbp 5
done
run "c2 default" --workload config2_tp53_chr17 --algo 3 --check-oracle-mbp 20
run "c3 default" --workload config3_16S_500Mbp --algo 3 --check-oracle-mbp 10
run "c5 L18 k0" --workload config5_sweep_1Gbp --primer-len 18 --k 0 --check-oracle-mbp 5
run "c5 L35 k4" --workload config5_sweep_1Gbp --primer-len 35 --k 4 --check-oracle-mbp 5
python - <<'PY'
import json
for l in open('gpurun_out/jit_sweep.jsonl'):
    d = json.loads(l)
    r = d['r']
    if 'failed' in r:
        print(d['tag'], 'FAILED'); continue
    print('%-22s algo %d scan %.3f ms main %.3f ms dg %.3f ms wall %.3f ms cand %d hits %d oracle %s' % (
        d['tag'], r['algo'], r['ms_scan_kernels'], r['ms_main_kernel'], r['ms_dg_kernel'], r['ms_find_wall_best'],
        r['candidates'], r['hits'], r.get('oracle_equal')))
PY
