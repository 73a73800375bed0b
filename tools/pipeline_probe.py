#!/usr/bin/env python3
"""How many calls in flight / host threads pay on a workload: wall time per step of the fused call
for (lanes, threads) combinations.  python tools/pipeline_probe.py --workload northstar_256fold_3Gbp"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', required=True)
    ap.add_argument('--scale', type=float, default=1.0)
    ap.add_argument('--steps', type=int, default=12)
    args = ap.parse_args()
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import synthetic
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    wl = synthetic.make_workload(args.workload, scale=args.scale)
    primers, k = wl['primers'], wl['k']
    var = [unambiguous_sequences(p) for p in primers]
    tdf.PCR_P.set(COND)
    ctx = tdf._ctx()
    cond = tdf.conditions()
    tmpl = ctx.template_from_sequences([(r.ctypes.data, len(r)) for r in wl['records']])
    lanes = [ctx] + [tdf.new_context(0) for _ in range(3)]
    qs = [c.prepare(primers, k, var, jit=True) for c in lanes]
    for c, q in zip(lanes, qs):
        for _ in range(3):
            c.find(tmpl, q, cond=cond)
        c.set_timing(False)
    out = {}
    for n_lanes, n_threads in ((1, 1), (2, 1), (2, 2), (3, 1), (4, 2), (4, 4)):
        per = args.steps

        def work(my):
            F = len(my)
            for i in range(min(F - 1, per)):
                lanes[my[i % F]].find_begin(tmpl, qs[my[i % F]], cond)
            for i in range(per):
                if i + F - 1 < per:
                    j = (i + F - 1) % F
                    lanes[my[j]].find_begin(tmpl, qs[my[j]], cond)
                h, d, st, dst = lanes[my[i % F]].find_end()

        groups = [list(range(t, n_lanes, n_threads)) for t in range(n_threads)]
        best = None
        for _ in range(2):
            for c in lanes:
                c.synchronize()
            t0 = time.perf_counter()
            ths = [threading.Thread(target=work, args=(g,)) for g in groups[1:]]
            for th in ths:
                th.start()
            work(groups[0])
            for th in ths:
                th.join()
            for c in lanes:
                c.synchronize()
            dt = (time.perf_counter() - t0) / (per * n_threads) * 1e3
            best = dt if best is None else min(best, dt)
        out['lanes%d_threads%d' % (n_lanes, n_threads)] = round(best, 4)
    print(json.dumps({'workload': args.workload, 'ms_per_step': out}))


if __name__ == '__main__':
    main()
