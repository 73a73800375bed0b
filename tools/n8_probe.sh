#!/bin/bash
# N = 8: calls in flight x host threads, with the host-side breakdown of the pipelined region
run() {
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 8 --steps ${STEPS:-60} --warmup 5 --no-e2e --no-north-star "$@" 2>gpurun_out/n8_probe.err | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); h = d['run']['host_ms_per_step']
print('$* N=8 step %.3f wall %s seq %.3f  on-device call %.3f  scan %.3f dg %.3f | host begin %.3f collect %.3f end %.3f cores %s parity %s' % (d['ms_per_step'], d['run']['timing'].split('wall clock')[1], d['sequential']['ms_per_step'], d['sequential']['ms_call_on_device'], d['ms_scan_kernels'], d['ms_dg_kernels'], h['find_begin'], h['collect_gathered'], h['find_end_incl_wait'], d['run']['host_cores'], d['parity']['equal_all']))"
}
nproc
run --in-flight 6 --host-threads 3
run --in-flight 4 --host-threads 2
run --in-flight 3 --host-threads 3
run --in-flight 2 --host-threads 1
STEPS=20 run --in-flight 6 --host-threads 3
