#!/bin/bash
# letter-planes-first against shift-the-code-words in the shared-planes kernel (config 4)
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in 0 1 0 1; do
  DEGENPRIMER_B200_JIT_SHIFT_PLANES=$v python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-parity --no-e2e --no-north-star 2>gpurun_out/ab_$v.err | tail -1 | python -c "
import sys, json; d = json.loads(sys.stdin.read())
print('shift_planes=$v step %.3f  on-device %.3f  scan %.3f dg %.3f' % (d['ms_per_step'], d['sequential']['ms_call_on_device'], d['ms_scan_kernels'], d['ms_dg_kernels']))"
done
