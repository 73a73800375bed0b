#!/usr/bin/env python3
"""Generate tests/golden/reference_golden.json.gz by RUNNING THE REFERENCE ITSELF.

`oracle/refshim.py` imports the unmodified DegenPrimer sources from /root/reference
under Python 3 (syntax rewritten in memory, Python-2 semantics restored by an AST
pass, Biopython stand-ins — the complete list of rewrites is in its docstring).
This script calls the reference's own

  SearchEngine._find / _find_mp / find_matches        DP/SearchEngine.py:257-372
  SearchEngine.compile_duplexes / compile_duplexes_mp  DP/SearchEngine.py:126-187
  Duplex / Dimer / dimer_dG_corrected                  DP/SecStructures.py:58-324, DP/TD_Functions.py:124-206
  Primer.generate_components / calculate_Tms           DP/Primer.py:107-140

on seeded inputs (the sequences of the reference's own tests: Test_SearchEngine.py:77-78,
Test_Product.py:28-29, Test_Primer.py:30, Test_SecStructures.py:28-35, README.md:72-96) and
records what they return.  These fixtures pin the oracle (tests/test_reference_golden.py, CPU)
and the CUDA path (same file, -m gpu).  Only runs in the build container (needs /root/reference).

usage: python tests/golden/make_reference_golden.py
"""
import gzip
import hashlib
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
warnings.simplefilter('ignore', SyntaxWarning)

from oracle import refshim                       # noqa: E402
from tests.util import plant_sites, random_dna   # noqa: E402

PRIMERS = ['ATATTCTACRACGGCTATCC', 'GAASGCRAAKATYGGGAAC', 'ATARTCTYCGAMGGCTATCC', 'NAAGGYTTAGAKGCGGAAG']
BIG = 'ATARTCTYCGAMGGCNATKCAGGNCTGRGGA'                      # Test_Primer.py:30, 512 variants, L=31
LONG35 = 'GAASGCRAAKATYGGGAACTTCAGGBCTGAGGAWC'                # L=35: windows span three 32-base words
SHORT18 = 'GAAGGCAAAGATCGGGAA'                                # L=18 (config 5's short end)
CONDS = {
    'readme': dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, DMSO=0.0, PCR_T=62.0),
    'test_ipcr': dict(Na=50e-3, Mg=3e-3, dNTP=300e-6, DNA=1e-10, DMSO=0.0, PCR_T=53.0),
    'default': dict(Na=50e-3, Mg=1.5e-3, dNTP=0.1e-3, DNA=5e-9, DMSO=0.0, PCR_T=60.0),
    'dmso_hot': dict(Na=35e-3, Mg=2.5e-3, dNTP=0.8e-3, DNA=2e-9, DMSO=5.0, PCR_T=68.0),
    'cold': dict(Na=100e-3, Mg=4e-3, dNTP=0.2e-3, DNA=1e-8, DMSO=1.5, PCR_T=45.0),
}
ALL_IUPAC = ('ACGTRYSWKMBDHVN', [0.22] * 4 + [0.12 / 11] * 11)
WITH_N = ('ACGTN', [0.24, 0.24, 0.24, 0.24, 0.04])
FEW_N = ('ACGTN', [0.2495] * 4 + [0.002])


def make_template(seed, n, alphabet, p, masters, k_plant, n_plant):
    rng = np.random.default_rng(seed)
    t = random_dna(rng, n, alphabet, p)
    if n_plant:
        t, _ = plant_sites(rng, t, masters, k_plant, n_plant)
    return t


def mask_of(matches):
    m = 0
    for i in matches:
        m |= 1 << int(i)
    return m


def duplex_digest(d):
    """sha1 over the strings a consumer sees: both sequences, print_most_stable() and str()."""
    text = '\x00'.join((d.fwd_seq, d.rev_seq, d.print_most_stable(), d.print_most_stable(False), str(d)))
    return hashlib.sha1(text.encode()).hexdigest()[:16]


def dump_duplex(d):
    """[name, fwd_3_overhang, have_3_matches, digest, [[fwd_matches as a bit mask, offset, dG, K, fwd_mismatch], ...]]"""
    return [d.name, d.fwd_3_overhang, d.have_3_matches, duplex_digest(d),
            [[mask_of(x.fwd_matches), x.offset, x.dG, x.K, x.fwd_mismatch] for x in d.dimers]]


def dump_results(res):
    if res is None:
        return None
    return [None if lst is None else [[int(pos), [dump_duplex(d) for d in dups]] for pos, dups in lst]
            for lst in res]


def main():
    ns = refshim.load()
    from multiprocessing import Event
    SE, WC, tdf = ns.SearchEngine.SearchEngine, ns.WorkCounter.WorkCounter, ns.tdf
    eng = SE(Event())

    def primer(master, conc=1e-6, pid='P'):
        return ns.Primer.Primer(ns.SeqRecord(ns.Seq(master), id=pid), conc, True)

    def record(t):
        return ns.SeqRecord(ns.Seq(t), id='T')

    gold = {'reference_commit': 'c610551c', 'generator': 'tests/golden/make_reference_golden.py via oracle/refshim.py',
            'conds': CONDS, 'scan': [], 'scan_mp': [], 'duplex': [], 'tm': [], 'dimer': [], 'errors': []}

    # ---- scan, serial _find (DP/SearchEngine.py:327-359)
    scan_specs = [
        (1, 6000, 1, ('ACGT', None), PRIMERS, 6),
        (2, 5000, 3, ('ACGT', None), PRIMERS, 6),
        (3, 4000, 6, WITH_N, PRIMERS[:2], 6),
        (4, 3000, 2, ALL_IUPAC, PRIMERS[1:3], 6),
        (5, 300, 25, ('ACGT', None), PRIMERS[:1], 6),          # k >= L: the max(1, L-k) clamp
        (6, 7000, 4, FEW_N, [BIG, LONG35, SHORT18], 8),
        (7, 2500, 0, ('ACGT', None), [LONG35, SHORT18, PRIMERS[3]], 5),
        (8, 4099, 4, ALL_IUPAC, [SHORT18, PRIMERS[1]], 6),     # one base more than 4096+3: ragged last chunk
        (9, 21, 1, ('ACGT', None), PRIMERS[:1], 0),            # T = L+1
        (10, 19, 2, ('ACGT', None), PRIMERS[1:2], 0),          # T = L (not a power of two)
        (11, 8192, 3, ('AC', [0.5, 0.5]), ['ACACACMCACACACACAC', 'NACACACASNCACACACW'], 0),   # low complexity: dense hits
    ]
    for seed, n, k, (alphabet, p), masters, n_plant in scan_specs:
        print("scan", seed, flush=True)
        t = make_template(seed, n, alphabet, p, masters, min(k, 4), n_plant)
        case = {'seed': seed, 'template': t, 'k': k, 'primers': {}}
        for m in masters:
            r = eng._find(WC(), record(t), primer(m), len(t), len(m), k)
            assert r[0] == t and r[3] == len(t) and r[4] == len(m)
            case['primers'][m] = {'fwd': r[5].tolist(), 'rev': r[6].tolist()}
        case['rev_seq_sha1'] = hashlib.sha1(r[1].encode()).hexdigest()
        gold['scan'].append(case)

    # ---- scan, _find_mp and the find_matches dispatch (DP/SearchEngine.py:241-324,362-372); templates by seed
    for seed, n, k, (alphabet, p), masters in ((21, 100003, 2, FEW_N, PRIMERS[:2]),
                                               (22, 262144 + 17, 3, ('ACGT', None), [PRIMERS[1], LONG35]),
                                               (23, 150001, 1, WITH_N, [SHORT18])):
        t = make_template(seed, n, alphabet, p, masters, k, 12)
        case = {'seed': seed, 'n': n, 'alphabet': alphabet, 'p': p, 'masters': masters, 'k_plant': k, 'n_plant': 12,
                'template_sha1': hashlib.sha1(t.encode()).hexdigest(), 'k': k, 'primers': {}}
        for m in masters:
            print('scan_mp', seed, m, flush=True)
            pr = primer(m)
            r_mp = None
            for _ in range(5):                  # the reference can start a finished job twice
                try:
                    r_mp = eng._find_mp(WC(), record(t), pr, len(t), len(m), k)
                    break
                except AssertionError:
                    continue
            r_se = eng._find(WC(), record(t), pr, len(t), len(m), k)
            assert r_mp is not None
            assert r_mp[5].tolist() == r_se[5].tolist() and r_mp[6].tolist() == r_se[6].tolist()
            case['primers'][m] = {'fwd': r_mp[5].tolist(), 'rev': r_mp[6].tolist()}
        gold['scan_mp'].append(case)

    # ---- duplex stage (DP/SearchEngine.py:106-187)
    for cname, master, k, seed, n, n_plant in (('readme', PRIMERS[1], 2, 11, 5000, 10),
                                               ('test_ipcr', PRIMERS[0], 3, 12, 5000, 10),
                                               ('default', PRIMERS[3], 2, 13, 5000, 10),
                                               ('dmso_hot', PRIMERS[2], 4, 14, 6000, 14),
                                               ('cold', LONG35, 4, 15, 6000, 10),
                                               ('test_ipcr', BIG, 3, 16, 3000, 3),
                                               ('cold', SHORT18, 4, 17, 30000, 12),
                                               # ambiguous template letters INSIDE the annealing windows: the NN
                                               # parameters of those stacks are averages over the expansions
                                               # (UnifiedNN.int_dPar_37, DP/UnifiedNN.py:224-246)
                                               ('readme', PRIMERS[1], 3, 18, 6000, 24),
                                               ('test_ipcr', PRIMERS[0], 4, 19, 6000, 24)):
        print('duplex', cname, master, flush=True)
        tdf.PCR_P.set(CONDS[cname])
        t = make_template(seed, n, *FEW_N, [master], k, n_plant)
        if seed in (18, 19):
            rng = np.random.default_rng(seed + 100)
            chars = list(t)
            for at in rng.choice(len(chars), size=len(chars) // 25, replace=False):
                chars[at] = 'NRYSWKMBDHV'[int(rng.integers(11))]
            t = ''.join(chars)
        pr = primer(master)
        r = eng._find(WC(), record(t), pr, len(t), len(master), k)
        mp = None
        for _ in range(5):
            try:
                mp = eng.compile_duplexes_mp(WC(), *r)
                break
            except AssertionError:
                continue
        gold['duplex'].append({'cond': cname, 'master': master, 'k': k, 'seed': seed, 'template': t,
                               'fwd': r[5].tolist(), 'rev': r[6].tolist(),
                               'mp': dump_results(mp),
                               'serial': dump_results(eng.compile_duplexes(WC(), *r))})

    # ---- per-variant Tm (DP/TD_Functions.py:60-101, DP/Primer.py:121-140)
    for cname, master, conc in (('readme', BIG, 1.5e-6), ('default', PRIMERS[1], 0.9e-6),
                                ('test_ipcr', PRIMERS[0], 0.9e-6), ('dmso_hot', LONG35, 0.25e-6),
                                ('cold', SHORT18, 2e-6)):
        tdf.PCR_P.set(CONDS[cname])
        pr = primer(master, conc)
        gold['tm'].append({'cond': cname, 'master': master, 'concentration': conc,
                           'num_components': pr.num_components,
                           'ids': [r.id for r in pr.seq_records],
                           'variants': [str(s) for s in pr.sequences],
                           'tm': [tdf.primer_template_Tm(s, pr.concentration) for s in pr.sequences],
                           'Tm_min': pr.Tm_min, 'Tm_max': pr.Tm_max, 'Tm_mean': pr.Tm_mean})

    # ---- hand-built dimers incl. non-zero offsets = dangling ends (Test_SecStructures.py:28-35)
    Duplex, Dimer = ns.SecStructures.Duplex, ns.SecStructures.Dimer
    rc = ns.SecStructures.reverse_complement
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.15e-6, DNA=1e-9, DMSO=0.0, PCR_T=60.0)
    tdf.PCR_P.set(cond)
    fwd, rev = 'AGAGAACGCAAAGATCGGGAAC', 'CTTGCGTTTCTAACCCTTG'[::-1]
    idx = list(range(3, 15)) + list(range(16, 22))
    d = Duplex(fwd, rev, dimer=Dimer(idx, 3))
    gold['dimer'].append({'cond': cond, 'fwd': fwd, 'rev': rev, 'matches': idx, 'offset': 3,
                          'dimers': [[list(x.fwd_matches), x.dG, x.K] for x in d.dimers], 'text': str(d)})
    for cname in ('readme', 'dmso_hot'):
        tdf.PCR_P.set(CONDS[cname])
        for a, b in (('GAAGGCAAAGATCGGGAAC', 'GTTCCCGATCTTTGCCTTC'), ('ATATTCTACGACGGCTATCC', 'GGATAGCCGTCGTAGAATAT'),
                     ('GAAGGCAAAGATCGGGAAC', 'GTTCCCGATGTTTGCGTTC')):
            for off in (-4, -3, -1, 0, 1, 2, 5):
                dm = Dimer.from_sequences(a, rc(b), off)
                if dm:
                    gold['dimer'].append({'cond': CONDS[cname], 'fwd': a, 'rev': b, 'matches': list(dm.fwd_matches),
                                          'offset': off, 'dG': tdf.dimer_dG_corrected(dm, a, b),
                                          'alternatives': [list(x.fwd_matches) for x in dm.alternatives]})

    # ---- error / dispatch behaviour (DP/SearchEngine.py:232-245,362-372)
    for t_len, p_len in ((10, 20), (20, 0), (20, 20), (19, 20)):
        try:
            SE._check_length_inequality(t_len, p_len)
            gold['errors'].append({'t_len': t_len, 'p_len': p_len, 'raises': None})
        except ValueError as e:
            gold['errors'].append({'t_len': t_len, 'p_len': p_len, 'raises': 'ValueError', 'message': str(e)})
    tdf.PCR_P.set(CONDS['default'])
    r = eng._find(WC(), record('ACGT' * 50), primer(PRIMERS[0]), 200, 20, 1)
    gold['no_hits'] = {'fwd': r[5].tolist(), 'rev': r[6].tolist(),
                       'compile_duplexes': eng.compile_duplexes(WC(), *r),
                       'compile_duplexes_mp': eng.compile_duplexes_mp(WC(), *r)}

    out = os.path.join(HERE, 'reference_golden.json.gz')
    with open(out, 'wb') as raw, gzip.GzipFile(fileobj=raw, mode='wb', mtime=0, filename='') as fh:
        fh.write(json.dumps(gold, separators=(',', ':')).encode())
    print('scan %d, scan_mp %d, duplex %d, tm %d, dimer %d -> %s (%d bytes)' % (
        len(gold['scan']), len(gold['scan_mp']), len(gold['duplex']), len(gold['tm']), len(gold['dimer']),
        out, os.path.getsize(out)))


if __name__ == '__main__':
    main()
