#!/usr/bin/env python3
"""Generate tests/golden/reference_dimers.json.gz by running the reference's own
SecStructures (self-dimers, cross-dimers, hairpins; DP/SecStructures.py:389-716) through
oracle/refshim.py — SURVEY §8(f) N3.  See make_reference_golden.py for how the shim works.

usage: python tests/golden/make_reference_dimers.py
"""
import gzip
import hashlib
import json
import os
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
warnings.simplefilter('ignore', SyntaxWarning)

from oracle import refshim                                         # noqa: E402
from tests.golden.make_reference_golden import CONDS, mask_of      # noqa: E402

# primers of the reference's tests (Test_SearchEngine.py:77-78, Test_Product.py:28-29, Test_Primer.py:30),
# plus sequences with palindromic cores / GC-rich 3' ends so that stable structures exist
MASTERS = ['ATATTCTACRACGGCTATCC', 'GAASGCRAAKATYGGGAAC', 'ATARTCTYCGAMGGCTATCC', 'NAAGGYTTAGAKGCGGAAG']
EXTRA = ['GGGAAATTTCCCGGGAAATTTCCC', 'ACGCGTACGCGTTAGCGCGC', 'GCGGCCGCTTTTGCGGCCGC', 'CCGGAATTCCGGAGCTCGAATTCGG',
         'AGAGAACGCAAAGATCGGGAAC', 'GTTCCCGATCTTTGCGTTCTCT', 'TTTTTTTTTTTTTTTTTT', 'GCGCGCGCGCGCGCGCGC',
         'GAAGGCAAAGATCGGGAACTTCAGGGCTGAGGAAC', 'CCTCAGCCCTGAAGTTCCCGA', 'ACGT']


def main():
    ns = refshim.load()
    SS, tdf = ns.SecStructures.SecStructures, ns.tdf
    unamb = sys.modules['DegenPrimer.SeqUtils'].unambiguous_sequences
    seqs = list(EXTRA)
    for m in MASTERS:
        v = unamb(m)
        seqs += [v[0], v[-1], v[len(v) // 2]]
    seqs = list(dict.fromkeys(seqs))

    def rec(s, i):
        return ns.SeqRecord(ns.Seq(s), id='s%d' % i)

    def dump_dimers(d):
        return [[mask_of(x.fwd_matches), x.offset, x.dG] for x in d.values()]

    gold = {'generator': 'tests/golden/make_reference_dimers.py via oracle/refshim.py', 'conds': {}, 'seqs': seqs,
            'cases': []}
    for cname in ('readme', 'dmso_hot', 'cold'):
        gold['conds'][cname] = CONDS[cname]
        tdf.PCR_P.set(CONDS[cname])
        for i, a in enumerate(seqs):
            s = SS(rec(a, i))
            case = {'cond': cname, 'a': i, 'b': i, 'dimers': dump_dimers(s._seq1_dimers),
                    'n_3prim': s._3prim_dimers, 'min_3prim_dG': s._min_3prim_dimer_dG,
                    'hairpins': [[list(h.fwd_matches), list(h.rev_matches), h.dG] for h in s._seq1_hairpins.values()],
                    'n_3prim_hairpins': s._3prim_hairpins}
            case['text_sha1'] = [hashlib.sha1(x.encode()).hexdigest() for x in (str(s), s.formatShort())]
            if i < 6 and cname == 'readme':
                case['text'] = [str(s), s.formatShort()]
            seeds = s._find_dimers(s._seq1, s._seq1)['hairpins']
            case['fronts'] = [[mask_of(x.fwd_matches), x.offset] for x in seeds.values()]
            gold['cases'].append(case)
            if cname != 'readme' and i % 3:
                continue
            for j, b in enumerate(seqs):
                if j == i or (cname != 'readme' and j % 2):
                    continue
                c = SS(rec(a, i), rec(b, j))
                gold['cases'].append({'cond': cname, 'a': i, 'b': j, 'dimers': dump_dimers(c._cross_dimers),
                                      'n_3prim': c._3prim_dimers, 'min_3prim_dG': c._min_3prim_dimer_dG,
                                      'text_sha1': [hashlib.sha1(x.encode()).hexdigest()
                                                    for x in (str(c), c.formatShort())]})
    out = os.path.join(HERE, 'reference_dimers.json.gz')
    with open(out, 'wb') as raw, gzip.GzipFile(fileobj=raw, mode='wb', mtime=0, filename='') as fh:
        fh.write(json.dumps(gold, separators=(',', ':'), default=lambda o: o.item()).encode())
    n_d = sum(len(c['dimers']) for c in gold['cases'])
    n_h = sum(len(c.get('hairpins', ())) for c in gold['cases'])
    print('%d sequences, %d pairs, %d stable dimers, %d hairpins -> %s (%d bytes)' % (
        len(seqs), len(gold['cases']), n_d, n_h, out, os.path.getsize(out)))


if __name__ == '__main__':
    main()
