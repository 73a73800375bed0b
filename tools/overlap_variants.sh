#!/bin/bash
# One GPU, 1/8 of the genome (a rank's share at N = 8, without the exchange): does a duplex / publish geometry that
# fits next to a resident scan block shorten the pipelined step?
for v in "" V2 V2p ""; do
  lib=""; [ -n "$v" ] && lib=$PWD/degenprimer_b200/libvariant_$v.so
  DEGENPRIMER_B200_LIB=$lib python bench.py --scale 0.125 --steps 60 --warmup 5 --no-cpu-baseline --no-parity --no-e2e 2>gpurun_out/ov_$v.err | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); n = d['north_star']
print('variant [$v] step %.3f  seq %.3f  on-device %.3f  scan %.3f dg %.3f | ns step %.3f scan %.3f dg %.3f' % (d['ms_per_step'], d['sequential']['ms_per_step'], d['sequential']['ms_call_on_device'], d['ms_scan_kernels'], d['ms_dg_kernels'], n['ms_per_step'], n['ms_scan_kernels'], n['ms_dg_kernels']))"
done
