#!/usr/bin/env python3
"""SURVEY §8(f) N3 measurement: all self- and cross-dimers of the V variants of a degenerate primer
(V*(V+1)/2 pairs, `SecStructures._find_dimers` each) in one dp_find_dimers launch, next to the CPU
oracle on a bounded sample of the same pairs.  Prints one JSON line.

usage: python tools/bench_dimers.py [--variants 256] [--reps 5] [--cpu-pairs 150]
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
    ap.add_argument('--master', default='ATARTCTYCGAMGGCNATKCAGGNCTGRGGA')     # Test_Primer.py:30, 512 variants
    ap.add_argument('--variants', type=int, default=256)
    ap.add_argument('--reps', type=int, default=5)
    ap.add_argument('--cpu-pairs', type=int, default=150)
    a = ap.parse_args()
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SecStructures import find_dimers_batch
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, DMSO=0.0, PCR_T=62.0)     # README.md:72-96
    tdf.PCR_P.set(cond)
    seqs = unambiguous_sequences(a.master)[:a.variants]
    V, L = len(seqs), len(seqs[0])
    pairs = np.array([(i, j) for i in range(V) for j in range(i, V)], dtype=np.uint32)
    ctx = tdf._ctx()
    c = tdf.conditions()
    ctx.find_dimers(seqs, pairs, c)                                               # warm-up
    ms_k, ms_call = [], []
    for _ in range(a.reps):
        t0 = time.perf_counter()
        rec, ms = ctx.find_dimers(seqs, pairs, c)
        ms_call.append((time.perf_counter() - t0) * 1e3)
        ms_k.append(ms)
    t0 = time.perf_counter()
    sets = find_dimers_batch(seqs, pairs)
    ms_objects = (time.perf_counter() - t0) * 1e3
    # hairpins of every variant: seeds = the self-dimers' full structures
    from degenprimer_b200.SecStructures import find_hairpins_batch
    self_sets = [sets[k] for k in range(len(pairs)) if pairs[k][0] == pairs[k][1]]
    find_hairpins_batch(self_sets)                                                # warm-up, table upload
    t0 = time.perf_counter()
    hsets, ms_hp = find_hairpins_batch(self_sets, return_ms=True)
    ms_hp_objects = (time.perf_counter() - t0) * 1e3
    n_seeds = sum(len(ds.fronts) for ds in self_sets)
    n_fronts = len(pairs) * (2 * L - 1)
    # candidates evaluated: n_runs*(n_runs+1)/2 per front
    full = rec['full_mask']
    runs = np.zeros(full.shape, dtype=np.int64)
    starts = full & ~(full << np.uint64(1))
    for b in range(64):
        runs += ((starts >> np.uint64(b)) & np.uint64(1)).astype(np.int64)
    n_eval = int((runs * (runs + 1) // 2).sum())
    # CPU oracle (Python restatement of the reference's loop) on a bounded sample of the same pairs
    from oracle import conditions as ocond
    from oracle.structures import find_dimers as o_find
    ocond.PCR_P.set(cond)
    rng = np.random.default_rng(0)
    sample = rng.choice(len(pairs), size=min(a.cpu_pairs, len(pairs)), replace=False)
    t0 = time.perf_counter()
    same = True
    for p in sample:
        i, j = pairs[p]
        dimers, _, n3, _ = o_find(seqs[i], seqs[j])
        ds = sets[p]
        same &= [(d.fwd_matches, d.offset, d.dG) for d in dimers] == [(d.fwd_matches, d.offset, d.dG) for d in ds.dimers]
        same &= n3 == ds.n_3prim_dimers
    cpu_s = time.perf_counter() - t0
    k = float(np.median(ms_k))
    print(json.dumps({
        'workload': 'all self+cross dimers of %d variants of %s (L=%d): %d pairs, %d fronts, %d dimer dG evaluations'
                    % (V, a.master, L, len(pairs), n_fronts, n_eval),
        'kernel': 'dp_find_dimers_kernel', 'ms_kernel': k, 'ms_call_host_to_host': float(np.median(ms_call)),
        'ms_with_python_objects': ms_objects,
        'pairs_per_s': len(pairs) / (k * 1e-3), 'dimer_dG_per_s': n_eval / (k * 1e-3),
        'stable_dimers': int(sum(len(s) for s in sets)),
        'hairpins': {'kernel': 'dp_find_hairpins_kernel', 'sequences': len(self_sets), 'seeds': n_seeds,
                     'ms_kernel': ms_hp, 'ms_with_python_objects': ms_hp_objects,
                     'stable_hairpins': int(sum(len(h) for h in hsets))},
        'cpu_oracle': {'pairs': int(len(sample)), 'seconds': cpu_s, 'pairs_per_s': len(sample) / cpu_s, 'cores': 1,
                       'kind': 'port', 'equal_to_gpu': bool(same)},
        'speedup_kernel_vs_cpu_core': (len(pairs) / (k * 1e-3)) / (len(sample) / cpu_s)}))


if __name__ == '__main__':
    main()
