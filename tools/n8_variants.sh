#!/bin/bash
# N = 8: library variants x calls in flight
run() {
  local v="$1"; shift
  local lib=""; [ -n "$v" ] && lib=$PWD/degenprimer_b200/libvariant_$v.so
  DEGENPRIMER_B200_LIB=$lib python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 8 --steps 30 --warmup 5 --no-e2e --no-north-star "$@" 2>gpurun_out/n8_$v.err | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); print('variant [$v] $* N=8 step %.3f  seq %.3f  on-device call %.3f  scan %.3f dg %.3f  parity %s' % (d['ms_per_step'], d['sequential']['ms_per_step'], d['sequential']['ms_call_on_device'], d['ms_scan_kernels'], d['ms_dg_kernels'], d['parity']['equal_all']))"
}
nproc
run E
run ""
run E --in-flight 6 --host-threads 3
