#!/bin/bash
# Closing call of the round: ncu --set full of one config-4 step and one north-star step (plain run first), the
# counters file refreshed from it, then the default bench line that reads those counters.
tag=${1:-r2j}
bash tools/ncu_full.sh ${tag}_full 21 6
ncu -i gpurun_out/${tag}_full.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
python tools/refresh_counters.py gpurun_out/${tag}_raw.csv "profiles/${tag}_ncu_full.txt (ncu --set full, round 2, closing kernels): smsp__inst_executed.sum, dram__bytes_read+write.sum of one launch on the whole 3.1 Gbp genome" && cp profiles/kernel_counters.json gpurun_out/${tag}_kernel_counters.json
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
n = d['north_star']; e = d['e2e']; i = d['int_roofline']
print('c4 %.3f ms/step scan %.3f dg %.3f | e2e %.2f ms | ns %.3f scan %.3f dg %.3f | parity %s | alu %.3f issued %.3f' % (
    d['ms_per_step'], d['ms_scan_kernels'], d['ms_dg_kernels'], e['ms_per_step'],
    n['ms_per_step'], n['ms_scan_kernels'], n['ms_dg_kernels'], d['parity']['equal_all'], i.get('frac_alu_pipe', 0), i.get('frac_issued', 0)))
PY
