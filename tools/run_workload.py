#!/usr/bin/env python3
"""Run one synthetic BASELINE workload through the fused device path and print a JSON line.

    python tools/run_workload.py --workload config4_20plex_3Gbp [--scale 0.1] [--steps 3]
                                 [--check-popc] [--check-oracle-mbp 20] [--k K] [--primer-len L]

Used for the per-config numbers under profiles/ (bench.py stays on BASELINE configs[1]).
--check-popc      : the per-window popcount kernel must give the same hit set (bit-exact).
--check-oracle-mbp: hits of the first N Mbp of record 0 against the C brute-force oracle.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', required=True)
    ap.add_argument('--scale', type=float, default=1.0)
    ap.add_argument('--steps', type=int, default=3)
    ap.add_argument('--k', type=int, default=None)
    ap.add_argument('--primer-len', type=int, default=None, help='config 5: prefix length of the 35-mer')
    ap.add_argument('--v16', action='store_true', help='config 5: the V=16 version of the 35-mer')
    ap.add_argument('--algo', type=int, default=0)
    ap.add_argument('--check-popc', action='store_true')
    ap.add_argument('--no-jit', action='store_true')
    ap.add_argument('--check-oracle-mbp', type=float, default=0.0)
    args = ap.parse_args()

    import torch
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import synthetic
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    t0 = time.perf_counter()
    wl = synthetic.make_workload(args.workload, scale=args.scale)
    t_gen = time.perf_counter() - t0
    primers, k = wl['primers'], wl['k']
    if args.k is not None:
        k = args.k
    if args.primer_len is not None:
        base = synthetic.SWEEP_35_V16 if args.v16 else synthetic.SWEEP_35
        primers = [base[:args.primer_len]]
    var = [unambiguous_sequences(p) for p in primers]
    n_var = sum(len(v) for v in var)
    tdf.PCR_P.set(COND)
    ctx = tdf._ctx()
    cond = tdf.conditions()
    srcs = [(r.ctypes.data, len(r)) for r in wl['records']]
    t0 = time.perf_counter()
    tmpl = ctx.template_from_sequences(srcs)
    t_up = time.perf_counter() - t0
    q = ctx.prepare(primers, k, var, jit=not args.no_jit)
    ctx.find(tmpl, q, cond=cond, algo=args.algo)
    ctx.find(tmpl, q, cond=cond, algo=args.algo)
    ms = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        hits, dimers, st, dst = ctx.find(tmpl, q, cond=cond, algo=args.algo)
        ms.append((time.perf_counter() - t0) * 1e3)
    best = min(ms)
    out = {
        'workload': args.workload, 'scale': args.scale, 'bp': wl['bp'], 'records': len(wl['records']),
        'primers': len(primers), 'primer_len': [len(p) for p in primers][:4], 'variants': n_var, 'k': k,
        'ms_find_wall_best': best, 'ms_find_wall_all': ms, 'ms_scan_kernels': st.ms_scan, 'ms_main_kernel': st.ms_main,
        'ms_dg_kernel': dst.ms_kernel, 'algo': st.algo, 'launches': st.launches + dst.launches,
        'windows': st.windows, 'candidates': st.candidates, 'hits': int(len(hits)), 'duplexes': dst.duplexes,
        'dimers': int(len(dimers)),
        'bp_variants_per_s': wl['bp'] * n_var / (best * 1e-3),
        'windows_per_s_main': st.windows / (st.ms_main * 1e-3) if st.ms_main else None,
        'hbm_gbs_main': 0.25 * wl['bp'] / (st.ms_main * 1e-3) / 1e9 if st.ms_main else None,
        's_generate': t_gen, 's_upload_pack': t_up, 'upload_gbs': wl['bp'] / t_up / 1e9,
    }
    if args.check_popc:
        h1, st1 = ctx.scan(tmpl, primers, k, algo=1)
        h2, st2 = ctx.scan(tmpl, primers, k, algo=2)
        out['popc_equal'] = bool(np.array_equal(h1, h2) and np.array_equal(h2, hits))
        out['ms_popc_kernel'] = st1.ms_main
        assert out['popc_equal'], 'bit-slice and popcount kernels disagree'
    if args.check_oracle_mbp > 0:
        from oracle import o2c
        n = int(min(len(wl['records'][0]), args.check_oracle_mbp * 1e6))
        text = wl['records'][0][:n].tobytes()
        sub = ctx.template_from_sequences([text])
        hs, _ = ctx.scan(sub, primers, k)
        ok = True
        t0 = time.perf_counter()
        for i, m in enumerate(primers):
            f, r = o2c.find(text, m, k)
            h = hs[hs['primer'] == i]
            ok &= bool(np.array_equal(h[h['strand'] == 0]['pos'].astype(np.int64), f) and
                       np.array_equal(h[h['strand'] == 1]['pos'].astype(np.int64), r))
        out['oracle_equal'] = ok
        out['oracle_bp'] = n
        out['s_oracle'] = time.perf_counter() - t0
        assert ok, 'GPU hits differ from the C oracle'
    print(json.dumps(out))


if __name__ == '__main__':
    main()
