#!/usr/bin/env python3
"""SURVEY §8(f) N4 measurement: bringing a template into HBM from a FASTA file (parse + H2D + pack),
from host ASCII already in memory (H2D + pack), and from the packed cache (dp_template_load).
Prints one JSON line.

usage: python tools/bench_packed.py [--mbp 83.257441] [--reps 3] [--dir /tmp]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--mbp', type=float, default=83.257441)
    ap.add_argument('--reps', type=int, default=3)
    ap.add_argument('--dir', default='/tmp')
    a = ap.parse_args()
    from degenprimer_b200 import _native as nat
    n = int(a.mbp * 1e6)
    rng = np.random.default_rng(2)
    seq = np.frombuffer(b'ACGT', dtype=np.uint8)[rng.integers(0, 4, size=n, dtype=np.uint8)]
    seq[n // 3:n // 3 + 3000000] = ord('N')                       # one long run like config 2
    fa, cache = os.path.join(a.dir, 'bench_packed.fasta'), os.path.join(a.dir, 'bench_packed.dpk')
    with open(fa, 'wb') as fh:
        fh.write(b'>chr17_like synthetic\n')
        lines = np.empty((n // 60, 61), dtype=np.uint8)
        lines[:, :60] = seq[:n // 60 * 60].reshape(-1, 60)
        lines[:, 60] = 10
        fh.write(lines.tobytes())
        fh.write(seq[n // 60 * 60:].tobytes() + b'\n')
    ctx = nat.context()
    blob = seq.tobytes()

    def timed(fn):
        best = None
        for _ in range(a.reps):
            t0 = time.perf_counter()
            t = fn()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
            last = t
            if _ < a.reps - 1:
                t.close()
        return best * 1e3, last

    ms_fasta, t_fa = timed(lambda: ctx.template_from_fasta(fa))
    t0 = time.perf_counter()
    t_fa.save(cache)
    ms_save = (time.perf_counter() - t0) * 1e3
    ms_ascii, t_as = timed(lambda: ctx.template_from_sequences([blob]))
    ms_load, t_ld = timed(lambda: ctx.template_load(cache))
    same = (t_ld.total_bases, t_ld.n_exception_runs, t_ld.lengths) == (t_fa.total_bases, t_fa.n_exception_runs, t_fa.lengths)
    masters, k = ['GAASGCRAAKATYGGGAAC', 'ATATTCTACRACGGCTATCC'], 1
    h_fa, _ = ctx.scan(t_fa, masters, k)
    h_ld, _ = ctx.scan(t_ld, masters, k)
    same = same and h_fa.tobytes() == h_ld.tobytes()
    same = same and t_ld.unpack(0, n // 3 - 50, 200) == t_fa.unpack(0, n // 3 - 50, 200)
    print(json.dumps({
        'workload': '%d bp, one record, 3 Mbp run of N' % n,
        'fasta_bytes': os.path.getsize(fa), 'packed_bytes': os.path.getsize(cache),
        'ms_from_fasta_file': ms_fasta, 'ms_from_host_ascii': ms_ascii, 'ms_from_packed_cache': ms_load,
        'ms_save_cache': ms_save, 'speedup_vs_fasta': ms_fasta / ms_load, 'speedup_vs_host_ascii': ms_ascii / ms_load,
        'packed_GBps': os.path.getsize(cache) / ms_load / 1e6, 'identical_state_and_hits': bool(same),
        'note': 'best of %d; files in the page cache' % a.reps}))
    for f in (fa, cache):
        os.remove(f)


if __name__ == '__main__':
    main()
