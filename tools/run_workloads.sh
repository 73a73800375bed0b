#!/bin/bash
# all BASELINE configs + the north-star pair through the fused device path -> gpurun_out/r1d_workloads.jsonl
out=gpurun_out/r1d_workloads.jsonl
: > $out
run() { timeout 900 python tools/run_workload.py "$@" 2>gpurun_out/wl.err | tail -1 >> $out || echo "{\"failed\": \"$*\"}" >> $out; }
run --workload config1_tp53_5Mbp --check-popc --check-oracle-mbp 5
run --workload config2_tp53_chr17 --check-popc --check-oracle-mbp 20
run --workload config3_16S_500Mbp --check-popc --check-oracle-mbp 10
run --workload config3_16S_256var_500Mbp --check-oracle-mbp 10
run --workload config5_sweep_1Gbp --primer-len 18 --k 0 --check-oracle-mbp 5
run --workload config5_sweep_1Gbp --primer-len 35 --k 4 --check-oracle-mbp 5
run --workload northstar_256fold_3Gbp --check-oracle-mbp 5
run --workload config4_20plex_3Gbp --check-oracle-mbp 2
python - <<'PY'
import json
for l in open('gpurun_out/r1d_workloads.jsonl'):
    d = json.loads(l)
    if 'failed' in d:
        print(d); continue
    print(d['workload'], d.get('primer_len'), 'k', d['k'], 'algo', d['algo'], 'wall %.3f scan %.3f dg %.3f' % (d['ms_find_wall_best'], d['ms_scan_kernels'], d['ms_dg_kernel']),
          {k: v for k, v in d.items() if k in ('popc_equal', 'oracle_equal', 'hits', 'dimers')})
PY
