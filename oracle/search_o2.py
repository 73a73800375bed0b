"""ORACLE O2 (test infrastructure, see oracle/README.md) — independent brute
force for the annealing-site scan.

Integer model of what DegenPrimer/SearchEngine.py:190-211 computes through
FFTs: score(q) = sum_j w(P[j], t[q+j]) / 3 with the thirds table below
(SURVEY §8a row A2), hit iff 3*score >= 3*max(1, L-k) - 1 (the float test
`score >= max(1, L-k) - 0.5` of :321-323/:355-357 on a value that is an exact
multiple of 1/3).  No FFT, no chunks, no slices: used to cross-check O1.
"""
import numpy as np

from .sequtils import EXPANSIONS, reverse_complement

BASES = 'ACGT'


def _weights():
    """w3[letter][base]: 3 = full match, 2 = partial, 0 = mismatch.

    Derived by hand from the root-of-unity maps (SearchEngine.py:44-79):
    letters whose map values are single roots (A C G T R Y K M, and the
    'major' base of B D H V) score a full 1; letters mapped to w1+w2 (= -1)
    on a plane score 2/3 for both bases of that plane.
    """
    w = {}
    for b in BASES:
        w[b] = {x: (3 if x == b else 0) for x in BASES}
    for l in 'RYKM':
        w[l] = {x: (3 if x in EXPANSIONS[l] else 0) for x in BASES}
    w['S'] = {'A': 0, 'C': 2, 'G': 2, 'T': 0}
    w['W'] = {'A': 2, 'C': 0, 'G': 0, 'T': 2}
    w['N'] = {'A': 2, 'C': 2, 'G': 2, 'T': 2}
    w['B'] = {'A': 0, 'C': 2, 'G': 2, 'T': 3}
    w['D'] = {'A': 2, 'C': 0, 'G': 3, 'T': 2}
    w['H'] = {'A': 2, 'C': 3, 'G': 0, 'T': 2}
    w['V'] = {'A': 3, 'C': 2, 'G': 2, 'T': 0}
    return w


W3 = _weights()
OTHER = 1      # any template byte outside ACGT scores 1/3 against every letter


def threshold3(p_len, mismatches):
    return 3 * max(1, p_len - mismatches) - 1


def score3(strand, master):
    """int array, one entry per window start 0..len(strand)-len(master)."""
    t = np.frombuffer(strand.encode('latin-1'), dtype=np.uint8)
    L = len(master)
    n = len(t) - L + 1
    if n <= 0:
        return np.zeros(0, dtype=np.int64)
    lut = np.full((L, 256), OTHER, dtype=np.int64)
    for j, letter in enumerate(master):
        for b in BASES:
            lut[j, ord(b)] = W3[letter][b]
    acc = np.zeros(n, dtype=np.int64)
    for j in range(L):
        acc += lut[j][t[j:j + n]]
    return acc


def find(fwd_seq, master, mismatches):
    """Same return convention as search_o1.find_serial()."""
    t_len, p_len = len(fwd_seq), len(master)
    if t_len < p_len or p_len == 0:
        raise ValueError('template shorter than primer, or empty primer')
    rev_seq = reverse_complement(fwd_seq)
    thr = threshold3(p_len, mismatches)
    fwd_matches = np.where(score3(fwd_seq, master) >= thr)[0]
    rev_matches = np.where(score3(rev_seq, master) >= thr)[0]
    return fwd_seq, rev_seq, t_len, p_len, fwd_matches, rev_matches
