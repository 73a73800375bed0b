"""ORACLE O1 (test infrastructure, see oracle/README.md) — literal restatement
of DegenPrimer/SearchEngine.py: cube-roots-of-unity maps (:36-79), pattern map
(:92-102), chunked scipy.fftpack correlation (:190-211), chunk sizing
(:214-229), serial `_find` (:328-359), sliced `_find_mp` (:280-325, slices run
in-process or in a multiprocessing pool), `compile_duplexes` (:160-187) and
`compile_duplexes_mp` ordering (:129-157).  This is also the timed CPU baseline.
"""
import multiprocessing
import os

import numpy as np
from scipy.fftpack import fft, ifft

from . import conditions as cond
from .sequtils import reverse_complement
from .structures import Duplex

W0 = 1
W1 = -1 / 2.0 + np.sqrt(3) / 2.0j     # SearchEngine.py:37-39 (sic: sqrt(3)/2j)
W2 = -1 / 2.0 - np.sqrt(3) / 2.0j

# template byte -> root of unity (SearchEngine.py:41-45)
T_AT = ((ord('A'), W1), (ord('T'), W2))
T_GC = ((ord('G'), W1), (ord('C'), W2))

# pattern letter -> (AT-map value, GC-map value)   (SearchEngine.py:47-79)
P_MAP = {
    'A': (W2, W0), 'T': (W1, W0), 'G': (W0, W2), 'C': (W0, W1),
    'R': (W2, W2), 'Y': (W1, W1), 'S': (W0, W2 + W1), 'W': (W2 + W1, W0),
    'K': (W1, W2), 'M': (W2, W1), 'B': (W1, W2 + W1), 'D': (W2 + W1, W2),
    'H': (W2 + W1, W1), 'V': (W2, W2 + W1), 'N': (W2 + W1, W2 + W1),
}

MAX_CHUNK = 2 ** 12                    # SearchEngine.py:83
CPU_COUNT = os.cpu_count() or 1


def map_pattern(pattern, map_len):
    """SearchEngine.py:92-102."""
    at = np.zeros(map_len, dtype=complex)
    gc = np.zeros(map_len, dtype=complex)
    for i, letter in enumerate(pattern):
        at[i], gc[i] = P_MAP[letter]
    return at, gc


def find_in_chunk(t_chunk, p_fft, correction, c_size, c_stride):
    """SearchEngine.py:190-211.  t_chunk: bytes."""
    chunk = np.frombuffer(t_chunk, dtype=np.int8).astype(complex)
    chunk = np.concatenate((chunk, np.zeros(c_size - len(chunk), dtype=complex)))
    t_at = np.zeros(len(chunk), dtype=complex)
    t_gc = np.zeros(len(chunk), dtype=complex)
    for k, v in T_AT:
        t_at[chunk == k] = v
    for k, v in T_GC:
        t_gc[chunk == k] = v
    at_score = ifft(fft(t_at[::-1]) * p_fft[0])[::-1][:c_stride]
    gc_score = ifft(fft(t_gc[::-1]) * p_fft[1])[::-1][:c_stride]
    score = at_score.real + gc_score.real
    return score * (2.0 / 3.0) + correction


def calculate_chunk_size(t_len, p_len, max_chunk=MAX_CHUNK):
    """SearchEngine.py:214-229 (Python-2 integer division kept)."""
    def rem(c):
        return (t_len // (c - p_len)) * (c - p_len) + c - t_len
    if t_len <= max_chunk:
        chunk = 2 ** int(np.ceil(np.log2(t_len)))
        if chunk % t_len == 0:
            return chunk
    else:
        chunk = max_chunk
    r = rem(chunk)
    min_chunk = 2 ** int(np.ceil(np.log2(2 * p_len)))
    max_rem = chunk // 2 + 1
    while r > max_rem and chunk > min_chunk:
        chunk //= 2
        r = rem(chunk)
    return max(chunk, min_chunk)


def check_length_inequality(t_len, p_len):
    """SearchEngine.py:232-238."""
    if t_len < p_len or p_len == 0:
        raise ValueError('SearchEngine._find: template sequence should be longer or equal to primer '
                         'sequence and both should be grater than zero.')


def mp_better(t_len):
    """SearchEngine.py:241-245."""
    return CPU_COUNT > 1 and t_len > 25000


def optimal_slices(t_len, p_len):
    """SearchEngine.py:248-254."""
    linear = max(CPU_COUNT, int(t_len * 1.75e-5 + 1.75))
    return min(60, linear, t_len // p_len)


def _threshold(p_len, mismatches):
    return max(1, p_len - mismatches) - 0.5      # SearchEngine.py:321,355


def find_serial(fwd_seq, master, mismatches):
    """SearchEngine.py:328-359 without the counter/abort plumbing.

    fwd_seq, master: str.  Returns (fwd_seq, rev_seq, t_len, p_len, fwd_matches,
    rev_matches); rev_matches index the reverse-complement string.
    """
    t_len, p_len = len(fwd_seq), len(master)
    check_length_inequality(t_len, p_len)
    c_size = calculate_chunk_size(t_len, p_len)
    c_stride = c_size - p_len
    maps = map_pattern(master, c_size)
    p_fft = (fft(maps[0]), fft(maps[1]))
    rev_seq = reverse_complement(fwd_seq)
    fwd_b = fwd_seq.encode('latin-1')
    rev_b = rev_seq.encode('latin-1')
    correction = np.full(c_stride, p_len / 3.0)
    fwd_score = np.zeros(t_len + c_stride)
    rev_score = np.zeros(t_len + c_stride)
    pos = 0
    while pos < t_len:
        front = min(t_len, pos + c_size)
        fwd_score[pos:pos + c_stride] = find_in_chunk(fwd_b[pos:front], p_fft, correction, c_size, c_stride)
        rev_score[pos:pos + c_stride] = find_in_chunk(rev_b[pos:front], p_fft, correction, c_size, c_stride)
        pos += c_stride
    thr = _threshold(p_len, mismatches)
    fwd_matches = np.where(fwd_score[:t_len - p_len + 1] >= thr)[0]
    rev_matches = np.where(rev_score[:t_len - p_len + 1] >= thr)[0]
    return fwd_seq, rev_seq, t_len, p_len, fwd_matches, rev_matches


def _find_in_slice(args):
    """SearchEngine.py:257-277."""
    start, end, fwd_b, rev_b, p_fft, correction, s_stride, c_size, c_stride = args
    fwd_score, rev_score = [], []
    pos = start
    while pos < end:
        front = min(end, pos + c_size)
        fwd_score.append(find_in_chunk(fwd_b[pos:front], p_fft, correction, c_size, c_stride))
        rev_score.append(find_in_chunk(rev_b[pos:front], p_fft, correction, c_size, c_stride))
        pos += c_stride
    return (start, np.concatenate(fwd_score)[:s_stride], np.concatenate(rev_score)[:s_stride])


_POOL_STATE = {}


def _slice_worker(job):
    start, end = job
    st = _POOL_STATE
    return _find_in_slice((start, end, st['fwd'], st['rev'], st['p_fft'], st['corr'],
                           st['s_stride'], st['c_size'], st['c_stride']))


def find_mp(fwd_seq, master, mismatches, workers=None):
    """SearchEngine.py:280-325.  The reference forks one process per slice
    (<= 60) and lets the OS schedule them; here the same slices run in a fork
    pool of `workers` processes (workers=0/1: in-process) — identical scores.
    """
    t_len, p_len = len(fwd_seq), len(master)
    check_length_inequality(t_len, p_len)
    s_size = t_len // optimal_slices(t_len, p_len) + p_len + 1
    s_stride = s_size - p_len
    c_size = calculate_chunk_size(s_size, p_len)
    c_stride = c_size - p_len
    maps = map_pattern(master, c_size)
    p_fft = (fft(maps[0]), fft(maps[1]))
    rev_seq = reverse_complement(fwd_seq)
    fwd_b = fwd_seq.encode('latin-1')
    rev_b = rev_seq.encode('latin-1')
    correction = np.full(c_stride, p_len / 3.0)
    jobs = []
    pos = 0
    while pos < t_len:
        jobs.append((pos, min(t_len, pos + s_size)))
        pos += s_stride
    scores_len = s_stride * len(jobs)
    scores = [np.zeros(scores_len), np.zeros(scores_len)]
    _POOL_STATE.update(fwd=fwd_b, rev=rev_b, p_fft=p_fft, corr=correction,
                       s_stride=s_stride, c_size=c_size, c_stride=c_stride)
    try:
        if workers and workers > 1:
            ctx = multiprocessing.get_context('fork')
            with ctx.Pool(workers) as pool:
                results = pool.map(_slice_worker, jobs, chunksize=1)
        else:
            results = [_slice_worker(j) for j in jobs]
    finally:
        _POOL_STATE.clear()
    for start, f, r in results:
        scores[0][start:start + len(f)] = f
        scores[1][start:start + len(r)] = r
    thr = _threshold(p_len, mismatches)
    fwd_matches = np.where(scores[0][:t_len - p_len + 1] >= thr)[0]
    rev_matches = np.where(scores[1][:t_len - p_len + 1] >= thr)[0]
    return fwd_seq, rev_seq, t_len, p_len, fwd_matches, rev_matches


def find_matches(fwd_seq, master, mismatches, workers=None):
    """SearchEngine.py:362-372: dispatch on mp_better."""
    check_length_inequality(len(fwd_seq), len(master))
    if mp_better(len(fwd_seq)):
        return find_mp(fwd_seq, master, mismatches, workers)
    return find_serial(fwd_seq, master, mismatches)


def compile_duplexes_for_position(position, strand_seq, primer, t_len, p_len, reverse):
    """SearchEngine.py:106-118: every variant at every site."""
    window = strand_seq[position:position + p_len]
    dups = []
    for var in primer.seq_records:
        d = Duplex(str(var.seq), window, name=var.id, revcomp=True)
        if d:
            dups.append(d)
    if reverse:
        return t_len + 1 - (position + p_len), tuple(dups)
    return position + p_len, tuple(dups)


def compile_duplexes(fwd_seq, rev_seq, primer, t_len, p_len, fwd_matches, rev_matches, mp_order=False):
    """SearchEngine.py:160-187 (serial; mp_order=False) or the result ordering
    of compile_duplexes_mp :129-157 (mp_order=True: both lists sorted by
    position, empty lists stay lists)."""
    if not len(fwd_matches) + len(rev_matches):
        return None
    with cond.AcquireParameters():
        fwd_results = []
        for pos in fwd_matches:
            r = compile_duplexes_for_position(int(pos), fwd_seq, primer, t_len, p_len, False)
            if r[1]:
                fwd_results.append(r)
        rev_results = []
        for pos in rev_matches:
            r = compile_duplexes_for_position(int(pos), rev_seq, primer, t_len, p_len, True)
            if r[1]:
                rev_results.append(r)
    if mp_order:
        if not fwd_results and not rev_results:
            return None
        fwd_results.sort(key=lambda x: x[0])
        rev_results.sort(key=lambda x: x[0])
        return fwd_results, rev_results
    if not fwd_results:
        fwd_results = None
    if not rev_results:
        rev_results = None
    if not fwd_results and not rev_results:
        return None
    return fwd_results, rev_results
