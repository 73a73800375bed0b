#!/usr/bin/env python3
"""Generate tests/golden/reference_mixture.json.gz by running the reference's own product pairing
(MixtureFactory.create_PCR_mixture / _sort_annealings, DP/PCR_Mixture.py:157-249; Product / Region,
DP/Product.py:30-215) on annealings its own SearchEngine produced — SURVEY §8(f) N1.  Everything goes
through oracle/refshim.py (see make_reference_golden.py).

usage: python tests/golden/make_reference_mixture.py
"""
import gzip
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
warnings.simplefilter('ignore', SyntaxWarning)

from oracle import refshim                                                # noqa: E402
from tests.golden.make_reference_golden import CONDS                      # noqa: E402
from tests.util import plant_sites, random_dna                            # noqa: E402

CASES = [
    # cond, primers [(id, master)], k, seed, template length, planted sites per primer, min/max amplicon,
    # with_exonuclease, include_side_annealings
    ('test_ipcr', [('F', 'ATATTCTACRACGGCTATCC'), ('R', 'GAASGCRAAKATYGGGAAC')], 2, 31, 6000, 14, 50, 2000, False, False),
    ('test_ipcr', [('F', 'ATATTCTACRACGGCTATCC'), ('R', 'GAASGCRAAKATYGGGAAC')], 2, 31, 6000, 14, 50, 2000, True, True),
    ('readme', [('F', 'ATARTCTYCGAMGGCTATCC'), ('R', 'NAAGGYTTAGAKGCGGAAG')], 3, 32, 9000, 20, 100, 1500, False, True),
    ('cold', [('A', 'GAAGGCAAAGATCGGGAA'), ('B', 'ATATTCTACGACGGCTATCC'), ('C', 'GAASGCRAAKATYGGGAAC')], 4, 33, 8000, 16, 30,
     4000, False, False),
    ('dmso_hot', [('F', 'ATATTCTACRACGGCTATCC')], 1, 34, 4000, 24, 20, 3000, False, True),      # one primer, both roles
    ('default', [('F', 'ATATTCTACRACGGCTATCC'), ('R', 'GAASGCRAAKATYGGGAAC')], 1, 35, 3000, 2, 500, 600, False, True),
]


class StubDuplex(object):
    """What Product reads of a Duplex: name, fwd_len, print_most_stable(), repr."""
    def __init__(self, name, n):
        self.name, self.fwd_len = name, n

    def print_most_stable(self, include_fwd_3_mismatch=True):
        return 'duplex %s %s\n' % (self.name, include_fwd_3_mismatch)

    def __repr__(self):
        return 'rep %s' % self.name


def main():
    ns = refshim.load()
    import DegenPrimer.PCR_Mixture as pm
    from multiprocessing import Event
    SE, WC, tdf = ns.SearchEngine.SearchEngine, ns.WorkCounter.WorkCounter, ns.tdf
    eng = SE(Event())
    gold = {'generator': 'tests/golden/make_reference_mixture.py via oracle/refshim.py', 'conds': CONDS, 'cases': []}
    for cond, primers, k, seed, n, n_plant, amin, amax, exo, side in CASES:
        tdf.PCR_P.set(CONDS[cond])
        rng = np.random.default_rng(seed)
        t = random_dna(rng, n, 'ACGTN', [0.2495] * 4 + [0.002])
        t, _ = plant_sites(rng, t, [m for _, m in primers], min(k, 2), n_plant)
        prs = [ns.Primer.Primer(ns.SeqRecord(ns.Seq(m), id=pid), 1e-6, True) for pid, m in primers]
        fwd, rev = [], []
        for p in prs:
            r = eng._find(WC(), ns.SeqRecord(ns.Seq(t), id='T'), p, len(t), len(p), k)
            ann = eng.compile_duplexes(WC(), *r)
            if ann is None:
                continue
            if ann[0]:
                fwd.extend(ann[0])
            if ann[1]:
                rev.extend(ann[1])
        n_sites = (len(fwd), len(rev))
        mf = pm.MixtureFactory(Event(), prs, amin, amax, with_exonuclease=exo, include_side_annealings=side)
        mix = mf.create_PCR_mixture(WC(), 'T synthetic', fwd, rev)
        case = {'cond': cond, 'primers': primers, 'k': k, 'seed': seed, 'template': t, 'min_amplicon': amin,
                'max_amplicon': amax, 'with_exonuclease': exo, 'include_side_annealings': side, 'n_sites': n_sites,
                'mixture': None}
        if mix is not None:
            def region(r):
                return [r.name, r.start, r.end, bool(r.forward)]
            case['mixture'] = {
                'reaction_id': mix.reaction_id,
                'products': sorted([p.start, p.end, p.fwd_ids, p.rev_ids, region(p.fwd_template), region(p.rev_template),
                                    len(p)] for p in mix.products.values()),
                'fwd_templates': [region(r) for r in mix.templates[1]],
                'rev_templates': [region(r) for r in mix.templates[0]],
                'annealings': [[region(tpl), [d.name for d in dups]] for dups, tpl in mix.annealings]}
        gold['cases'].append(case)
        print(cond, n_sites, None if mix is None else (len(mix.products), len(mix.annealings)), flush=True)
    # report text of Product / Region with stand-in duplexes (DP/Product.py:54-66,128-162)
    import DegenPrimer.Product as rp
    a = rp.Product('t seq', 101, 300, [StubDuplex('P_1', 20), StubDuplex('P_2', 25)], [StubDuplex('R_1', 18)])
    a.quantity, a.cycles = 2.5e-7, 12
    gold['texts'] = {'str': str(a), 'pretty': a.pretty_print(), 'pretty_no_name_no_3': a.pretty_print(False, False),
                     'region': str(rp.Region('x', 3, 9, False)), 'fwd_ids': a.fwd_ids,
                     'fwd_template': [a.fwd_template.start, a.fwd_template.end]}
    # the reference's own Product smoke test (Tests/DegenPrimer_Tests/Test_Product.py:27-39): two duplexes of
    # DEGENERATE primers with their reverse complements (NN parameters averaged over the expansions)
    tdf.PCR_P.set(CONDS['default'])
    rcomp, Duplex = ns.SecStructures.reverse_complement, ns.SecStructures.Duplex
    fp, rp_ = 'ATARTCTYCGAMGGCTATCC', 'NAAGGYTTAGAKGCGGAAG'
    d1, d2 = Duplex(fp, rcomp(fp)), Duplex(rp_, rcomp(rp_))
    tp = rp.Product('test_template', 50, 150, (d1,), (d2,))
    gold['test_product'] = {'cond': 'default', 'fwd': fp, 'rev': rp_, 'str': str(tp), 'pretty': tp.pretty_print(),
                            'dG': [d1.dG, d2.dG]}
    out = os.path.join(HERE, 'reference_mixture.json.gz')
    with open(out, 'wb') as raw, gzip.GzipFile(fileobj=raw, mode='wb', mtime=0, filename='') as fh:
        fh.write(json.dumps(gold, separators=(',', ':'), default=lambda o: o.item()).encode())
    print('%d cases -> %s (%d bytes)' % (len(gold['cases']), out, os.path.getsize(out)))


if __name__ == '__main__':
    main()
