#!/bin/bash
# exchange / duplex geometry variants at N = 2 (same exchange bytes per rank as N = 8): pipelined and sequential step
for v in "" C E F; do
  lib=""; [ -n "$v" ] && lib=$PWD/degenprimer_b200/libvariant_$v.so
  DEGENPRIMER_B200_LIB=$lib python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-north-star 2>gpurun_out/n2_$v.err | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); print('variant [$v] N=2 step %.3f  seq %.3f  on-device call %.3f  scan %.3f dg %.3f  parity %s' % (d['ms_per_step'], d['sequential']['ms_per_step'], d['sequential']['ms_call_on_device'], d['ms_scan_kernels'], d['ms_dg_kernels'], d['parity']['equal_all']))"
done
