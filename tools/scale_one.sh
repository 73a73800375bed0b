#!/bin/bash
# one driver-shaped bench line at N GPUs -> gpurun_out/<tag>_scale_n<N>.json     bash tools/scale_one.sh <tag> <N>
tag=${1:-r2i}; n=${2:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
  bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${tag}_scale_n$n.json 2> gpurun_out/${tag}_scale_n$n.err
echo "N=$n rc=$?"; tail -2 gpurun_out/${tag}_scale_n$n.err | cut -c1-300
python - <<PY
import json
d = json.load(open('gpurun_out/${tag}_scale_n$n.json'))
ns = d.get('north_star', {})
print('N=%d  %.3f ms/step  %.3e  seq %.3f ms  on-device %.3f scan %.3f dg %.3f  e2e %.2f ms  parity %s  | north-star %.3f ms | host %s' % (
    $n, d['ms_per_step'], d['value'], d['sequential']['ms_per_step'], d['sequential']['ms_call_on_device'], d['ms_scan_kernels'], d['ms_dg_kernels'],
    d.get('e2e', {}).get('ms_per_step', 0), d['parity']['equal_all'] if d.get('parity') else None,
    ns.get('ms_per_step', 0), d['run']['host_ms_per_step']))
PY
