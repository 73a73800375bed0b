#!/bin/bash
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
for cfg in "100 1" "50 1" "0 1" "100 0" "0 0"; do
  set -- $cfg
  DEGENPRIMER_B200_JIT_FMA=$1 DEGENPRIMER_B200_JIT_EXACT=$2 python bench.py --steps 200 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('fma $1 exact $2', 'ms/step %.4f'%d['ms_per_step'], 'main %.4f'%d['roofline']['ms_per_launch'], 'scan %.4f'%d['ms_scan_kernels'], 'frac %.3f'%d['roofline']['frac'])"
done
for W in config3_16S_500Mbp config4_20plex_3Gbp northstar_256fold_3Gbp "config5_sweep_1Gbp --primer-len 18 --k 0" "config5_sweep_1Gbp --primer-len 35 --k 4"; do
  for F in 100 0; do
  DEGENPRIMER_B200_JIT_FMA=$F python tools/run_workload.py --workload $W --check-oracle-mbp 2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('fma $F', d[\"workload\"], d[\"algo\"], d[\"launches\"], \"wall\", round(d[\"ms_find_wall_best\"],3), \"scan\", round(d[\"ms_scan_kernels\"],3), \"dg\", round(d[\"ms_dg_kernel\"],3), d.get(\"oracle_equal\"))"
  done
done
