"""The oracle (and, under -m gpu, the CUDA path) against outputs of THE REFERENCE ITSELF.

tests/golden/reference_golden.json.gz was recorded by tests/golden/make_reference_golden.py, which
runs the unmodified DegenPrimer sources (SearchEngine._find/_find_mp, compile_duplexes[_mp], Duplex,
Dimer, dimer_dG_corrected, Primer.calculate_Tms) through oracle/refshim.py.  These tests need neither
/root/reference nor the shim.  Hit positions, dimer structures and strings must be identical; dG, K
and Tm are compared with `==` (bar: 0.01 kcal/mol, 0.05 C).
"""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import conditions as ocond
from oracle import o2c
from oracle import search_o1 as o1
from oracle import search_o2 as o2
from oracle import thermo
from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord
from oracle.sequtils import reverse_complement
from oracle.structures import Dimer as ODimer, Duplex as ODuplex

from tests.util import plant_sites, random_dna

with gzip.open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_golden.json.gz'), 'rt') as _fh:
    REF = json.load(_fh)
CONDS = REF['conds']


def make_template(seed, n, alphabet, p, masters, k_plant, n_plant):
    rng = np.random.default_rng(seed)
    t = random_dna(rng, n, alphabet, p)
    if n_plant:
        t, _ = plant_sites(rng, t, masters, k_plant, n_plant)
    return t


def mp_template(case):
    t = make_template(case['seed'], case['n'], case['alphabet'], case['p'], case['masters'],
                      case['k_plant'], case['n_plant'])
    assert hashlib.sha1(t.encode()).hexdigest() == case['template_sha1']
    return t


def mask_of(matches):
    m = 0
    for i in matches:
        m |= 1 << int(i)
    return m


def duplex_digest(d):
    text = '\x00'.join((d.fwd_seq, d.rev_seq, d.print_most_stable(), d.print_most_stable(False), str(d)))
    return hashlib.sha1(text.encode()).hexdigest()[:16]


def check_results(got, want, exact=True):
    """got: (fwd_results, rev_results) of Duplex objects; want: the recorded compact form."""
    assert (got is None) == (want is None)
    if got is None:
        return 0
    n = 0
    for g_list, w_list in zip(got, want):
        assert (g_list is None) == (w_list is None)
        if g_list is None:
            continue
        assert [int(p) for p, _ in g_list] == [p for p, _ in w_list]
        for (_, gd), (_, wd) in zip(g_list, w_list):
            assert len(gd) == len(wd)
            for d, (name, overhang, have3, digest, dimers) in zip(gd, wd):
                assert d.name == name and d.fwd_3_overhang == overhang and d.have_3_matches == have3
                assert len(d.dimers) == len(dimers)
                for x, (mask, offset, dG, K, fwd_mismatch) in zip(d.dimers, dimers):
                    assert mask_of(x.fwd_matches) == mask and x.offset == offset
                    assert abs(x.dG - dG) <= 0.01                   # the bar
                    if exact:
                        assert x.dG == dG and x.K == K              # what is achieved
                    assert x.fwd_mismatch == fwd_mismatch
                assert duplex_digest(d) == digest
                n += 1
    return n


# ------------------------------------------------------------------ oracle vs reference (CPU)
def test_oracle_scan_equals_reference():
    for case in REF['scan']:
        t = case['template']
        assert hashlib.sha1(reverse_complement(t).encode()).hexdigest() == case['rev_seq_sha1']
        for m, want in case['primers'].items():
            r = o1.find_serial(t, m, case['k'])
            assert r[4].tolist() == want['fwd'] and r[5].tolist() == want['rev']
            r2 = o2.find(t, m, case['k'])
            assert r2[4].tolist() == want['fwd'] and r2[5].tolist() == want['rev']
            f, r3 = o2c.find(t, m, case['k'])
            assert f.tolist() == want['fwd'] and r3.tolist() == want['rev']


def test_oracle_find_mp_equals_reference():
    for case in REF['scan_mp']:
        t = mp_template(case)
        for m, want in case['primers'].items():
            r = o1.find_matches(t, m, case['k'])            # dispatches to the sliced path (T > 25000)
            assert r[4].tolist() == want['fwd'] and r[5].tolist() == want['rev']
            f, rv = o2c.find(t, m, case['k'])
            assert f.tolist() == want['fwd'] and rv.tolist() == want['rev']


def test_oracle_duplexes_equal_reference():
    total = 0
    for case in REF['duplex']:
        ocond.PCR_P.set(CONDS[case['cond']])
        p = OPrimer(OSeqRecord(case['master'], id='P'), 1e-6)
        r = o1.find_serial(case['template'], case['master'], case['k'])
        assert r[4].tolist() == case['fwd'] and r[5].tolist() == case['rev']
        for key, mp in (('mp', True), ('serial', False)):
            got = o1.compile_duplexes(r[0], r[1], p, r[2], r[3], r[4], r[5], mp_order=mp)
            total += check_results(got, case[key])
    assert total > 5000
    nh = REF['no_hits']
    assert nh['fwd'] == [] and nh['rev'] == [] and nh['compile_duplexes'] is None and nh['compile_duplexes_mp'] is None
    r = o1.find_serial('ACGT' * 50, 'ATATTCTACRACGGCTATCC', 1)
    assert len(r[4]) == 0 and len(r[5]) == 0
    assert o1.compile_duplexes(r[0], r[1], OPrimer(OSeqRecord('ATATTCTACRACGGCTATCC', id='P'), 1e-6),
                               r[2], r[3], r[4], r[5]) is None


def test_oracle_tm_equals_reference():
    for case in REF['tm']:
        ocond.PCR_P.set(CONDS[case['cond']])
        p = OPrimer(OSeqRecord(case['master'], id='P'), case['concentration'])
        assert p.calculate_Tms()
        assert [r.id for r in p.seq_records] == case['ids'] and p.str_sequences == case['variants']
        assert len(p.seq_records) == case['num_components']
        assert [thermo.primer_template_Tm(s, p.concentration) for s in p.str_sequences] == case['tm']
        assert (p.Tm_min, p.Tm_max, p.Tm_mean) == (case['Tm_min'], case['Tm_max'], case['Tm_mean'])


def test_oracle_dimers_equal_reference():
    for case in REF['dimer']:
        ocond.PCR_P.set(case['cond'])
        if 'dimers' in case:                                # Test_SecStructures.py:28-35
            d = ODuplex(case['fwd'], case['rev'], dimer=ODimer(case['matches'], case['offset']))
            assert [[list(x.fwd_matches), x.dG, x.K] for x in d.dimers] == case['dimers']
            assert str(d) == case['text']
        else:
            dm = ODimer.from_sequences(case['fwd'], reverse_complement(case['rev']), case['offset'])
            assert list(dm.fwd_matches) == case['matches']
            assert [list(x.fwd_matches) for x in dm.alternatives] == case['alternatives']
            assert thermo.dimer_dG_corrected(dm, case['fwd'], case['rev']) == case['dG']


def test_oracle_errors_equal_reference():
    for case in REF['errors']:
        if case['raises']:
            with pytest.raises(ValueError) as e:
                o1.check_length_inequality(case['t_len'], case['p_len'])
            assert str(e.value) == case['message']
        else:
            o1.check_length_inequality(case['t_len'], case['p_len'])


# ------------------------------------------------------------------ CUDA path vs reference (-m gpu)
def _gpu_primer(master, conc=1e-6):
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    gp = Primer(SeqRecord(Seq(master), id='P'), conc)
    assert gp.generate_components()
    return gp


def _gpu_record(t):
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    return SeqRecord(Seq(t), id='T', description='T synthetic')


def _gpu_engine():
    import threading
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.WorkCounter import WorkCounter
    return SearchEngine(threading.Event()), WorkCounter


@pytest.mark.gpu
@pytest.mark.parametrize('idx', range(len(REF['scan'])))
def test_gpu_scan_equals_reference(idx):
    case = REF['scan'][idx]
    se, WC = _gpu_engine()
    rec = _gpu_record(case['template'])
    for m, want in case['primers'].items():
        r = se.find_matches(WC(), rec, _gpu_primer(m), case['k'])
        assert r[3] == len(case['template']) and r[4] == len(m)
        assert r[5].tolist() == want['fwd'] and r[6].tolist() == want['rev']


@pytest.mark.gpu
@pytest.mark.parametrize('idx', range(len(REF['scan_mp'])))
def test_gpu_scan_long_equals_reference(idx):
    case = REF['scan_mp'][idx]
    se, WC = _gpu_engine()
    rec = _gpu_record(mp_template(case))
    for m, want in case['primers'].items():
        r = se.find_matches(WC(), rec, _gpu_primer(m), case['k'])
        assert r[5].tolist() == want['fwd'] and r[6].tolist() == want['rev']


@pytest.mark.gpu
@pytest.mark.parametrize('idx', range(len(REF['duplex'])))
def test_gpu_duplexes_equal_reference(idx):
    from degenprimer_b200 import TD_Functions as tdf
    case = REF['duplex'][idx]
    tdf.PCR_P.set(CONDS[case['cond']])
    se, WC = _gpu_engine()
    gp = _gpu_primer(case['master'])
    m = se.find_matches(WC(), _gpu_record(case['template']), gp, case['k'])
    assert m[5].tolist() == case['fwd'] and m[6].tolist() == case['rev']
    assert check_results(se.compile_duplexes_mp(WC(), *m), case['mp']) > 0
    assert check_results(se.compile_duplexes(WC(), *m), case['serial']) > 0


@pytest.mark.gpu
def test_gpu_no_hits_and_errors_equal_reference():
    se, WC = _gpu_engine()
    gp = _gpu_primer('ATATTCTACRACGGCTATCC')
    m = se.find_matches(WC(), _gpu_record('ACGT' * 50), gp, 1)
    assert m[5].tolist() == REF['no_hits']['fwd'] and m[6].tolist() == REF['no_hits']['rev']
    assert se.compile_duplexes(WC(), *m) is None and se.compile_duplexes_mp(WC(), *m) is None
    for case in REF['errors']:
        if case['raises']:
            with pytest.raises(ValueError) as e:
                se._check_length_inequality(case['t_len'], case['p_len'])
            assert str(e.value) == case['message']
        else:
            se._check_length_inequality(case['t_len'], case['p_len'])
    with pytest.raises(ValueError):
        se.find_matches(WC(), _gpu_record('ACGTACGTAC'), gp, 1)


@pytest.mark.gpu
@pytest.mark.parametrize('idx', range(len(REF['tm'])))
def test_gpu_tm_equals_reference(idx):
    from degenprimer_b200 import TD_Functions as tdf
    case = REF['tm'][idx]
    tdf.PCR_P.set(CONDS[case['cond']])
    gp = _gpu_primer(case['master'], case['concentration'])
    assert gp.calculate_Tms()
    assert gp.num_components == case['num_components']
    assert [r.id for r in gp.seq_records] == case['ids'] and gp.str_sequences == case['variants']
    tms = np.asarray(tdf.primer_template_Tms(gp.sequences, gp.concentration), dtype=np.float64)
    assert np.abs(tms - np.asarray(case['tm'])).max() <= 0.05          # the bar
    assert np.abs(tms - np.asarray(case['tm'])).max() < 1e-9           # what is achieved
    for got, want in ((gp.Tm_min, case['Tm_min']), (gp.Tm_max, case['Tm_max']), (gp.Tm_mean, case['Tm_mean'])):
        assert abs(got - want) < 1e-9


@pytest.mark.gpu
def test_gpu_dimers_equal_reference():
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SecStructures import Duplex, Dimer
    from degenprimer_b200.SecStructures import reverse_complement as g_rc
    n = 0
    for case in REF['dimer']:
        tdf.PCR_P.set(case['cond'])
        if 'dimers' in case:                                # Test_SecStructures.py:28-35
            d = Duplex(case['fwd'], case['rev'], dimer=Dimer(case['matches'], case['offset']))
            assert [list(x.fwd_matches) for x in d.dimers] == [x[0] for x in case['dimers']]
            for x, (_, dG, K) in zip(d.dimers, case['dimers']):
                assert abs(x.dG - dG) <= 0.01 and abs(x.dG - dG) < 1e-9 and abs(x.K - K) <= 1e-9 * K
            assert str(d) == case['text']
        else:
            dm = Dimer.from_sequences(case['fwd'], g_rc(case['rev']), case['offset'])
            assert list(dm.fwd_matches) == case['matches']
            assert [list(x.fwd_matches) for x in dm.alternatives] == case['alternatives']
            got = tdf.dimer_dG_corrected(dm, case['fwd'], case['rev'])
            assert abs(got - case['dG']) <= 0.01 and abs(got - case['dG']) < 1e-9
        n += 1
    assert n == len(REF['dimer']) and n > 30


# ------------------------------------------------------------------ host-side mirrors vs reference (CPU, no kernels)
def test_host_mirrors_equal_reference():
    """Primer component generation / ids and Dimer construction are host code of the facade."""
    from degenprimer_b200.SecStructures import Dimer
    from degenprimer_b200.SecStructures import reverse_complement as g_rc
    for case in REF['tm']:
        gp = _gpu_primer(case['master'], case['concentration'])
        assert gp.num_components == case['num_components']
        assert [r.id for r in gp.seq_records] == case['ids'] and gp.str_sequences == case['variants']
    for case in REF['dimer']:
        if 'dimers' in case:
            continue
        dm = Dimer.from_sequences(case['fwd'], g_rc(case['rev']), case['offset'])
        assert list(dm.fwd_matches) == case['matches'] and dm.offset == case['offset']
        assert [list(x.fwd_matches) for x in dm.alternatives] == case['alternatives']
    for case in REF['scan']:
        assert hashlib.sha1(g_rc(case['template']).encode()).hexdigest() == case['rev_seq_sha1']


def test_device_tables_reproduce_reference_dG_on_ambiguous_windows():
    """The flattened NN tables the kernels index (UnifiedNN.device_tables: stack / rstack averaged over the
    expansions of an ambiguous template letter, loop_dS, terminal terms) walked in Python exactly as
    dp_dimer_dg_device walks them (offset 0: left-aligned primer/window duplexes) give the reference's dG for
    every dimer of the duplex cases whose windows contain N / IUPAC letters — no GPU involved."""
    from degenprimer_b200 import _native as nat
    from degenprimer_b200.UnifiedNN import UnifiedNN
    T = UnifiedNN.device_tables()
    letter = {c: i for i, c in enumerate(nat.LETTERS)}
    comp = [2, 3, 0, 1, 5, 4, 6, 7, 9, 8, 13, 12, 11, 10, 14]

    def stack(tab, x, y, w):
        return tab[(x * 4 + y) * 15 + w]

    def dimer_dg(s, r, mask, PCR_T, Na_eq_log):
        L = len(s)
        fm = [i for i in range(L) if mask >> i & 1]
        f0, f1 = fm[0], fm[-1]
        dG, dH, dS = 0.0, T.ini_dH, T.ini_dS
        if f0 == 0:
            if s[0] in (0, 2):
                dH += T.ter_dH
                dS += T.ter_dS
        else:
            dG += T.terminal_mismatch
        if f1 < L - 1:
            dG += T.terminal_mismatch
        elif s[L - 1] in (0, 2):
            dH += T.ter_dH
            dS += T.ter_dS
        for a, b in zip(fm, fm[1:]):
            gap = b - a
            if gap < 4:
                assert r[a] == comp[s[a]]
                dH += stack(T.stack_dH, s[a], s[a + 1], r[a + 1])
                dS += stack(T.stack_dS, s[a], s[a + 1], r[a + 1])
                if gap > 1:
                    dH += stack(T.rstack_dH, s[b - 1], s[b], r[b - 1])
                    dS += stack(T.rstack_dS, s[b - 1], s[b], r[b - 1])
            else:
                dS += T.loop_dS[gap - 1]
        dS += T.dS_Na * (f1 - f0) * Na_eq_log
        return dG + dH - (PCR_T - T.K0) * dS / 1000.0

    checked = ambiguous = 0
    for case in REF['duplex'][-2:]:
        ocond.PCR_P.set(CONDS[case['cond']])
        p = OPrimer(OSeqRecord(case['master'], id='P'), 1e-6)
        variants = dict(zip([r.id for r in p.seq_records], p.str_sequences))
        t, L = case['template'], len(case['master'])
        rc = reverse_complement(t)
        for strand, lst in enumerate(case['serial']):
            for pos, dups in lst or []:
                window = t[pos - L:pos] if strand == 0 else rc[len(t) + 1 - pos - L:len(t) + 1 - pos]
                ambiguous += any(c not in 'ACGT' for c in window)
                r = [comp[letter[c]] for c in window]                     # 3'->5' complement strand under the primer
                for name, _, _, _, dimers in dups:
                    s = [letter[c] for c in variants[name]]
                    for mask, offset, dG, _, _ in dimers:
                        assert offset == 0
                        assert dimer_dg(s, r, mask, ocond.PCR_P.PCR_T, ocond.PCR_P.Na_eq_log) == dG
                        checked += 1
    assert checked > 300 and ambiguous >= 10
