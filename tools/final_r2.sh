#!/bin/bash
# Round-2 closing call: GPU suite, the default bench line, then the launch list of one short run
# (plain first, then the same command under ncu).   bash tools/final_r2.sh <tag>
tag=${1:-r2i}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --in-flight 1 --host-threads 1"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
echo "ncu rc=$?"
python - <<PY
import json
d = json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
n = d['north_star']; e = d['e2e']
print('c4 %.3f ms/step scan %.3f dg %.3f | e2e %.2f ms h2d %d | ns %.3f scan %.3f dg %.3f | parity %s | host %s' % (
    d['ms_per_step'], d['ms_scan_kernels'], d['ms_dg_kernels'], e['ms_per_step'], e['h2d_bytes_per_step'],
    n['ms_per_step'], n['ms_scan_kernels'], n['ms_dg_kernels'], d['parity']['equal_all'], d['run']['host_ms_per_step']))
print('cpu', d['cpu_baseline']['value'], 'value', d['value'], 'e2e', e['value'], 'roofline', d['roofline']['frac'], d['int_roofline'])
PY
