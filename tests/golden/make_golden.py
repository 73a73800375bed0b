#!/usr/bin/env python3
"""Generate tests/golden/*.json from the oracle (O1, the literal FFT restatement).

The reference ships no golden vectors for this path; these fixtures freeze the ORACLE
(the reference's own outputs are recorded by make_reference_golden.py next to this file):
they catch regressions in oracle/ and give the GPU tests fixed inputs with
recorded outputs.  Inputs reuse the sequences of the reference's own smoke
tests (Tests/DegenPrimer_Tests/Test_SearchEngine.py:77-78, Test_Product.py:28-29,
Test_SecStructures.py:28-35, Test_Primer.py:30, README.md:72-96).

usage: python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import conditions as ocond          # noqa: E402
from oracle import search_o1 as o1              # noqa: E402
from oracle import thermo                       # noqa: E402
from oracle.primer import Primer, SeqRecord     # noqa: E402
from oracle.structures import Dimer, Duplex     # noqa: E402
from tests.util import plant_sites, random_dna  # noqa: E402

PRIMERS = ['ATATTCTACRACGGCTATCC', 'GAASGCRAAKATYGGGAAC', 'ATARTCTYCGAMGGCTATCC', 'NAAGGYTTAGAKGCGGAAG']
CONDS = {
    'readme': dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, DMSO=0.0, PCR_T=62.0),
    'test_ipcr': dict(Na=50e-3, Mg=3e-3, dNTP=300e-6, DNA=1e-10, DMSO=0.0, PCR_T=53.0),
    'default': dict(Na=50e-3, Mg=1.5e-3, dNTP=0.1e-3, DNA=5e-9, DMSO=0.0, PCR_T=60.0),
}


def dump_results(res):
    if res is None:
        return None
    out = []
    for lst in res:
        if lst is None:
            out.append(None)
            continue
        out.append([[pos, [[d.name, [[list(dm.fwd_matches), dm.dG, dm.K, dm.fwd_mismatch] for dm in d.dimers]]
                           for d in dups]] for pos, dups in lst])
    return out


def scan_case(seed, n, k, alphabet, p, primers):
    rng = np.random.default_rng(seed)
    t = random_dna(rng, n, alphabet, p)
    t, _ = plant_sites(rng, t, primers, min(k, 4), 6)
    case = {'seed': seed, 'template': t, 'k': k, 'primers': {}}
    for m in primers:
        r = o1.find_serial(t, m, k)
        case['primers'][m] = {'fwd': r[4].tolist(), 'rev': r[5].tolist()}
    return case


def main():
    gold = {'scan': [], 'duplex': [], 'tm': [], 'dimer': []}
    gold['scan'].append(scan_case(1, 6000, 1, 'ACGT', None, PRIMERS))
    gold['scan'].append(scan_case(2, 5000, 3, 'ACGT', None, PRIMERS))
    gold['scan'].append(scan_case(3, 4000, 6, 'ACGTN', [0.24, 0.24, 0.24, 0.24, 0.04], PRIMERS[:2]))
    gold['scan'].append(scan_case(4, 3000, 2, 'ACGTRYSWKMBDHVN', [0.22] * 4 + [0.12 / 11] * 11, PRIMERS[1:3]))
    gold['scan'].append(scan_case(5, 300, 25, 'ACGT', None, PRIMERS[:1]))       # k >= L: the max(1, L-k) clamp
    for cname, master, k, seed in (('readme', PRIMERS[1], 2, 11), ('test_ipcr', PRIMERS[0], 3, 12),
                                   ('default', PRIMERS[3], 2, 13)):
        ocond.PCR_P.set(CONDS[cname])
        rng = np.random.default_rng(seed)
        t = random_dna(rng, 5000, 'ACGTN', [0.2495] * 4 + [0.002])
        t, _ = plant_sites(rng, t, [master], k, 10)
        p = Primer(SeqRecord(master, id='P'), 1e-6)
        r = o1.find_serial(t, master, k)
        gold['duplex'].append({
            'cond': cname, 'master': master, 'k': k, 'template': t,
            'mp': dump_results(o1.compile_duplexes(r[0], r[1], p, r[2], r[3], r[4], r[5], mp_order=True)),
            'serial': dump_results(o1.compile_duplexes(r[0], r[1], p, r[2], r[3], r[4], r[5], mp_order=False))})
    for cname, master, conc in (('readme', 'ATARTCTYCGAMGGCNATKCAGGNCTGRGGA', 1.5e-6),
                                ('default', PRIMERS[1], 0.9e-6), ('test_ipcr', PRIMERS[0], 0.9e-6)):
        ocond.PCR_P.set(CONDS[cname])
        p = Primer(SeqRecord(master, id='P'), conc)
        p.calculate_Tms()
        gold['tm'].append({'cond': cname, 'master': master, 'concentration': conc,
                           'ids': [r.id for r in p.seq_records][:4] + [p.seq_records[-1].id],
                           'tm': [thermo.primer_template_Tm(s, p.concentration) for s in p.str_sequences],
                           'Tm_min': p.Tm_min, 'Tm_max': p.Tm_max, 'Tm_mean': p.Tm_mean})
    # hand-built dimers incl. non-zero offsets (dangling ends) — Test_SecStructures.py:28-35
    ocond.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.15e-6, DNA=1e-9, DMSO=0.0, PCR_T=60.0))
    fwd, rev = 'AGAGAACGCAAAGATCGGGAAC', 'CTTGCGTTTCTAACCCTTG'[::-1]
    idx = list(range(3, 15)) + list(range(16, 22))
    d = Duplex(fwd, rev, dimer=Dimer(idx, 3))
    gold['dimer'].append({'cond': dict(Na=50e-3, Mg=3e-3, dNTP=0.15e-6, DNA=1e-9, DMSO=0.0, PCR_T=60.0),
                          'fwd': fwd, 'rev': rev, 'matches': idx, 'offset': 3,
                          'dimers': [[list(x.fwd_matches), x.dG, x.K] for x in d.dimers],
                          'text': str(d)})
    for off in (-3, -1, 0, 2):
        a, b = 'GAAGGCAAAGATCGGGAAC', 'GTTCCCGATCTTTGCCTTC'
        from oracle.sequtils import reverse_complement
        dm = Dimer.from_sequences(a, reverse_complement(b), off)
        if dm:
            gold['dimer'].append({'cond': None, 'fwd': a, 'rev': b, 'matches': list(dm.fwd_matches), 'offset': off,
                                  'dG': thermo.dimer_dG_corrected(dm, a, b)})
    with open(os.path.join(HERE, 'oracle_golden.json'), 'w') as fh:
        json.dump(gold, fh)
    print('scan cases %d, duplex %d, tm %d, dimer %d' % tuple(len(gold[k]) for k in ('scan', 'duplex', 'tm', 'dimer')))


if __name__ == '__main__':
    main()
