#!/bin/bash
# N = 8, 4, 2, 1 bench lines back to back on one 8-GPU box -> gpurun_out/<tag>_scale_n<N>.json
tag=${1:-r2c}
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${tag}_scale_n$n.json 2> gpurun_out/${tag}_scale_n$n.err
  echo "N=$n rc=$?"; tail -2 gpurun_out/${tag}_scale_n$n.err | cut -c1-300
done
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-parity > gpurun_out/${tag}_scale_n1.json 2> gpurun_out/${tag}_scale_n1.err
echo "N=1 rc=$?"
python - <<PY
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.load(open('gpurun_out/${tag}_scale_n%d.json' % n))
    except Exception as e:
        print(n, 'no line', e); continue
    base = base or d['value']
    ns = d.get('north_star', {})
    print('N=%d  %.3f ms/step  %.3e  x%.2f  seq %.3f ms  scan %.3f dg %.3f  e2e %.2f ms  parity %s  | north-star %.3f ms %.3e' % (
        n, d['ms_per_step'], d['value'], d['value'] / base, d['sequential']['ms_per_step'], d['ms_scan_kernels'], d['ms_dg_kernels'],
        d.get('e2e', {}).get('ms_per_step', 0), d['parity']['equal_all'] if d.get('parity') else None,
        ns.get('ms_per_step', 0), ns.get('value', 0)))
PY
