#!/bin/bash
# duplex-kernel geometry variants (LIST, warps per block, blocks per SM) on config 4 and the north-star pair
for v in "" B C D; do
  lib=""; [ -n "$v" ] && lib=$PWD/degenprimer_b200/libvariant_$v.so
  for wl in config4_20plex_3Gbp northstar_256fold_3Gbp; do
    DEGENPRIMER_B200_LIB=$lib python tools/run_workload.py --workload $wl --steps 5 2>/dev/null | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); print('variant [$v] %-24s dg %.3f ms  scan %.3f  wall %.3f  dimers %d' % (d['workload'], d['ms_dg_kernel'], d['ms_scan_kernels'], d['ms_find_wall_best'], d['dimers']))"
  done
done
