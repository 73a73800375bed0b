"""CPU tests of the oracle itself: O1 (literal FFT) == O2 (integer model) == C
brute force, the thirds table, hand-derived dG cases and the golden fixtures."""
import json
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import conditions as ocond
from oracle import o2c, search_o1 as o1, search_o2 as o2, thermo, unified_nn as nn
from oracle.primer import Primer, SeqRecord
from oracle.sequtils import reverse_complement, unambiguous_sequences
from oracle.structures import Dimer, Duplex

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'oracle_golden.json')))
IUPAC = 'ACGTRYSWKMBDHVN'


def test_expansion_order_and_ids():
    # earliest degenerate position varies fastest; N expands A,T,C,G (SeqUtils.py:30-60)
    assert unambiguous_sequences('ARNC')[:5] == ['AAAC', 'AGAC', 'AATC', 'AGTC', 'AACC']
    p = Primer(SeqRecord('ATARTCTYCGAMGGCNATKCAGGNCTGRGGA', id='x'), 1e-6)      # Test_Primer.py:30
    assert p.num_components == 512 and p.seq_records[0].id == 'x_001' and p.seq_records[-1].id == 'x_512'
    q = Primer(SeqRecord('ACGT', id='q'), 1e-6)
    assert q.num_components == 1 and q.seq_records[0].id == 'q'
    assert reverse_complement('acgtnRYKMbdhv') == 'BDHVKMRYNACGT'
    with pytest.raises(KeyError):
        reverse_complement('ACGU')


def test_thirds_table_is_what_the_fft_maps_compute():
    """Each (letter, base) weight through the literal root-of-unity maps."""
    for letter in IUPAC:
        for base in 'ACGTN':
            t = base + 'ACGTACGTAC'
            m = letter
            maps = o1.map_pattern(m, 16)
            from scipy.fftpack import fft
            sc = o1.find_in_chunk(t.encode(), (fft(maps[0]), fft(maps[1])), np.full(15, 1 / 3.0), 16, 15)[0]
            want = o2.W3[letter][base] if base in 'ACGT' else o2.OTHER
            assert abs(3 * sc - want) < 1e-9, (letter, base, sc)


@settings(max_examples=40, deadline=None)
@given(st.integers(0, 2 ** 31), st.integers(5, 30), st.integers(0, 35), st.sampled_from(['ACGT', 'ACGTN', IUPAC]))
def test_o1_equals_o2_property(seed, L, k, alphabet):
    rng = np.random.default_rng(seed)
    T = int(rng.integers(L, 6000))
    t = ''.join(rng.choice(list(alphabet), T))
    m = ''.join(rng.choice(list(IUPAC), L))
    a = o1.find_serial(t, m, k)
    b = o2.find(t, m, k)
    assert np.array_equal(a[4], b[4]) and np.array_equal(a[5], b[5])
    if T > 300:
        c = o1.find_mp(t, m, k)
        assert np.array_equal(c[4], b[4]) and np.array_equal(c[5], b[5])
    f, r = o2c.find(t, m, k)
    assert np.array_equal(f, b[4]) and np.array_equal(r, b[5])


def test_length_errors():
    with pytest.raises(ValueError):
        o1.find_serial('ACGT', 'ACGTACGT', 1)
    with pytest.raises(ValueError):
        o2.find('ACGT', '', 1)


def test_tables_complete():
    # every "WC pair next to anything" context exists after the reversed aliases
    comp = dict(zip('ACGT', 'TGCA'))
    for x in 'ACGT':
        for y in 'ACGT':
            for w in 'ACGT':
                assert (x + y + '/' + comp[x] + w) in nn.STACKS
    assert nn.INI.dH == 0.2 and nn.INI.dS == -5.7 and nn.TER.dH == 2.2 and nn.TER.dS == 6.9
    assert abs(nn.TERMINAL_MISMATCH + 0.72) < 1e-15
    assert nn.internal_loop_dG_37(3) == 3.2 + 2 * nn.TERMINAL_MISMATCH
    # averaged key: N on the template strand
    p = nn.int_dPar_37('AC', 'TN')
    want = sum(nn.STACKS['AC/T' + b].dH for b in 'ATCG') / 4.0
    assert abs(p.dH - want) < 1e-12


def test_hand_derived_dg():
    """Perfect 19-mer duplex, summed by hand from the table rows."""
    ocond.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0))
    p = 'GAAGGCAAAGATCGGGAAC'
    comp = dict(zip('ACGT', 'TGCA'))
    dH, dS = 0.2, -5.7                      # ini; no terminal AT on either end (G..C)
    for i in range(len(p) - 1):
        row = nn.STACKS[p[i:i + 2] + '/' + comp[p[i]] + comp[p[i + 1]]]
        dH += row.dH
        dS += row.dS
    dS += 0.368 * 18 * ocond.PCR_P.Na_eq_log
    want = dH - (62.0 + 273.15) * dS / 1000.0
    d = Duplex(p, p, revcomp=True)
    assert len(d.dimers) == 1 and abs(d.dG - want) < 1e-12
    assert abs(d.dG + 10.4234) < 1e-3 and abs(d.K / 6.27e6 - 1) < 1e-2        # SURVEY 8c smoke values
    w = p[:6] + 'T' + p[7:]                  # one interior mismatch
    d2 = Duplex(p, w, revcomp=True)
    assert [len(x.fwd_matches) for x in d2.dimers] == [18, 12]
    assert abs(d2.dimers[0].dG + 8.1858) < 1e-3 and abs(d2.dimers[1].dG + 6.0674) < 1e-3
    assert d2.dimers[0].fwd_mismatch is False


def test_alternatives_order():
    d = Dimer([0, 1, 2, 4, 5, 7, 8, 10, 11])
    got = [x.fwd_matches for x in d.alternatives]
    assert got == [(0, 1, 2, 4, 5, 7, 8), (0, 1, 2, 4, 5), (0, 1, 2), (4, 5, 7, 8, 10, 11), (4, 5, 7, 8),
                   (4, 5), (7, 8, 10, 11), (7, 8), (10, 11)]
    assert Dimer([3, 4, 5]).alternatives == [] and Dimer([3]).alternatives == []


def test_golden_scan():
    for case in GOLD['scan']:
        for m, want in case['primers'].items():
            r = o1.find_serial(case['template'], m, case['k'])
            assert r[4].tolist() == want['fwd'] and r[5].tolist() == want['rev']
            r2 = o2.find(case['template'], m, case['k'])
            assert r2[4].tolist() == want['fwd'] and r2[5].tolist() == want['rev']


def test_golden_duplex_and_tm():
    from tests.golden.make_golden import CONDS, dump_results
    for case in GOLD['duplex']:
        ocond.PCR_P.set(CONDS[case['cond']])
        p = Primer(SeqRecord(case['master'], id='P'), 1e-6)
        r = o1.find_serial(case['template'], case['master'], case['k'])
        for key, mp in (('mp', True), ('serial', False)):
            got = dump_results(o1.compile_duplexes(r[0], r[1], p, r[2], r[3], r[4], r[5], mp_order=mp))
            assert json.loads(json.dumps(got)) == case[key]
    for case in GOLD['tm']:
        ocond.PCR_P.set(CONDS[case['cond']])
        p = Primer(SeqRecord(case['master'], id='P'), case['concentration'])
        assert p.calculate_Tms()
        assert [thermo.primer_template_Tm(s, p.concentration) for s in p.str_sequences] == case['tm']
        assert (p.Tm_min, p.Tm_max, p.Tm_mean) == (case['Tm_min'], case['Tm_max'], case['Tm_mean'])
