#!/bin/bash
show() { python -c "
import json,sys; d=json.loads(open('$1').read()); print('$2', 'value %.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'seq %.4f'%d['sequential']['ms_per_step'], 'e2e ms %.3f'%d['e2e']['ms_per_step'], 'hits', d['hits_per_step'])"; }
for cfg in "2 1" "2 2" "4 2" "3 3" "4 4" "6 3"; do
  set -- $cfg
  python bench.py --no-cpu-baseline --in-flight $1 --host-threads $2 > gpurun_out/l.json 2> gpurun_out/l.err || tail -3 gpurun_out/l.err
  show gpurun_out/l.json "N=1 F=$1 T=$2"
done
for cfg in "2 1" "4 2" "4 4"; do
  set -- $cfg
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --in-flight $1 --host-threads $2 > gpurun_out/l2.json 2> gpurun_out/l2.err || tail -3 gpurun_out/l2.err
  show gpurun_out/l2.json "N=2 F=$1 T=$2"
done
