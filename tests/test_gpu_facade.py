"""SearchEngine facade vs the oracle: same signatures, same results (-m gpu)."""
import pickle
import threading

import numpy as np
import pytest

from oracle import conditions as ocond
from oracle import search_o1 as o1
from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord

from tests.util import random_dna, plant_sites

pytestmark = pytest.mark.gpu

COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.3e-3, DNA=1e-10, PCR_T=53.0)    # Test_iPCR.py:39-57


def _setup(master, k, n=40000, seed=3, alphabet='ACGT', p=None):
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    rng = np.random.default_rng(seed)
    t = random_dna(rng, n, alphabet, p)
    t, _ = plant_sites(rng, t, [master], min(k, 4), 16)
    ocond.PCR_P.set(COND)
    tdf.PCR_P.set(COND)
    op = OPrimer(OSeqRecord(master, id='P'), 0.9e-6)
    gp = Primer(SeqRecord(Seq(master), id='P'), 0.9e-6)
    gp.generate_components()
    return t, op, gp, SeqRecord(Seq(t), id='tmpl', description='tmpl synthetic')


def _same_results(got, want):
    assert (got is None) == (want is None)
    if got is None:
        return
    for g_list, w_list in zip(got, want):
        assert (g_list is None) == (w_list is None)
        if g_list is None:
            continue
        assert [p for p, _ in g_list] == [p for p, _ in w_list]
        for (_, gd), (_, wd) in zip(g_list, w_list):
            assert len(gd) == len(wd)
            for a, b in zip(gd, wd):
                assert a.name == b.name and a.fwd_seq == b.fwd_seq and a.rev_seq == b.rev_seq
                assert a.fwd_len == b.fwd_len and a.rev_len == b.rev_len
                assert a.fwd_3_overhang == b.fwd_3_overhang
                assert a.have_3_matches == b.have_3_matches
                assert len(a.dimers) == len(b.dimers)
                for x, y in zip(a.dimers, b.dimers):
                    assert x.fwd_matches == y.fwd_matches and x.offset == y.offset
                    assert abs(x.dG - y.dG) <= 0.01          # bar: 0.01 kcal/mol
                    assert x.dG == y.dG                      # achieved: bit-exact
                    assert x.K == y.K
                    assert x.fwd_mismatch == y.fwd_mismatch
                assert a.print_most_stable() == b.print_most_stable()
                assert str(a) == str(b)


@pytest.mark.parametrize('master,k', [('ATATTCTACRACGGCTATCC', 1), ('GAASGCRAAKATYGGGAAC', 3),
                                      ('NAAGGYTTAGAKGCGGAAG', 2)])
def test_find_matches_and_compile(master, k):
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.WorkCounter import WorkCounter
    t, op, gp, rec = _setup(master, k)
    se = SearchEngine(threading.Event())
    m = se.find_matches(WorkCounter(), rec, gp, k)
    ref = o1.find_matches(t, master, k)
    assert len(m) == 7 and m[2] is gp and m[3] == len(t) and m[4] == len(master)
    assert np.array_equal(m[5], ref[4]) and np.array_equal(m[6], ref[5])
    assert m[5].dtype == np.int64
    assert len(m[0]) == len(t) and m[0][10:50] == t[10:50] and m[1][5:45] == ref[1][5:45]
    want_mp = o1.compile_duplexes(ref[0], ref[1], op, ref[2], ref[3], ref[4], ref[5], mp_order=True)
    want_serial = o1.compile_duplexes(ref[0], ref[1], op, ref[2], ref[3], ref[4], ref[5], mp_order=False)
    _same_results(se.compile_duplexes_mp(WorkCounter(), *m), want_mp)
    _same_results(se.compile_duplexes(WorkCounter(), *m), want_serial)
    # plain strings in the 7-tuple (what PCR_Optimizer hands back) work too
    m2 = (ref[0], ref[1], gp, m[3], m[4], m[5], m[6])
    _same_results(se.compile_duplexes_mp(WorkCounter(), *m2), want_mp)


def test_find_returns_shelf_and_pickles():
    from degenprimer_b200.SearchEngine import SearchEngine, find
    from degenprimer_b200.WorkCounter import WorkCounter
    from degenprimer_b200.tmpStorage import from_shelf
    master, k = 'GAASGCRAAKATYGGGAAC', 2
    t, op, gp, rec = _setup(master, k, seed=9)
    ref = o1.find_matches(t, master, k)
    want = o1.compile_duplexes(ref[0], ref[1], op, ref[2], ref[3], ref[4], ref[5], mp_order=True)
    c = WorkCounter()
    path = find(c, rec, gp, k, threading.Event())
    assert isinstance(path, str)
    assert abs(c.percent() - 100.0) < 1e-9
    got = from_shelf(path)
    _same_results(got, want)
    assert isinstance(got[0][0][1], tuple)      # materialised on pickling
    pickle.loads(pickle.dumps(got))


def test_batch_find_and_errors(tmp_path):
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.SeqView import SeqView
    from degenprimer_b200.WorkCounter import WorkCounter
    from degenprimer_b200.tmpStorage import from_shelf
    master, k = 'ATATTCTACRACGGCTATCC', 1
    t, op, gp, rec = _setup(master, k, n=9000, seed=21)
    rng = np.random.default_rng(1)
    seqs = {'zeta': t, 'alpha': random_dna(rng, 500), 'mid': plant_sites(rng, random_dna(rng, 3000), [master], 1, 6)[0]}
    fa = tmp_path / 'templates.fasta'
    with open(fa, 'w') as fh:
        for name, s in seqs.items():
            fh.write('>%s synthetic %s\n' % (name, name))
            for i in range(0, len(s), 70):
                fh.write(s[i:i + 70].lower() + '\n')          # lower case: the loader folds it
    view = SeqView(upper=True)
    assert view.load(str(fa))
    assert view.keys() == ('alpha', 'mid', 'zeta') and len(view) == 3
    assert str(view['zeta'].seq) == t and len(view[1]) == 3000
    se = SearchEngine()
    res = se.batch_find(WorkCounter(), view.subview(view.keys()), gp, k, copy_data=True)
    for name, s in seqs.items():
        ref = o1.find_serial(s, master, k)
        want = o1.compile_duplexes(ref[0], ref[1], op, ref[2], ref[3], ref[4], ref[5], mp_order=False)
        if want is None:
            assert name not in res
        else:
            _same_results(from_shelf(res[name]), want)
    with pytest.raises(ValueError):
        se.find_matches(WorkCounter(), view['alpha'][:0] if False else _short(), gp, k)
    ev = threading.Event()
    ev.set()
    assert SearchEngine(ev).find_matches(WorkCounter(), rec, gp, k) is None
    assert SearchEngine(ev).find(WorkCounter(), rec, gp, k) is None


def _short():
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    return SeqRecord(Seq('ACGTACGT'), id='short')


def test_primer_tm_and_duplex_constructor():
    """Per-variant Tm (Primer.calculate_Tms) and a hand-built Duplex with a
    non-zero offset (Test_SecStructures.py:28-35: dangling-end branches)."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    from degenprimer_b200.SecStructures import Duplex, Dimer
    from oracle.structures import Duplex as ODuplex, Dimer as ODimer
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.15e-6, DNA=1e-9, PCR_T=60.0)
    ocond.PCR_P.set(cond)
    tdf.PCR_P.set(cond)
    gp = Primer(SeqRecord(Seq('ATARTCTYCGAMGGCNATKCAGGNCTGRGGA'), id='big'), 1.5e-6, generate_components=True)
    op = OPrimer(OSeqRecord('ATARTCTYCGAMGGCNATKCAGGNCTGRGGA', id='big'), 1.5e-6)
    assert op.calculate_Tms()
    assert gp.num_components == 512 and [r.id for r in gp.seq_records] == [r.id for r in op.seq_records]
    for a, b in ((gp.Tm_min, op.Tm_min), (gp.Tm_max, op.Tm_max), (gp.Tm_mean, op.Tm_mean)):
        assert abs(a - b) <= 0.05 and abs(a - b) < 1e-9
    idx = tuple(range(3, 15)) + tuple(range(16, 22))
    fwd, rev = 'AGAGAACGCAAAGATCGGGAAC', 'CTTGCGTTTCTAACCCTTG'[::-1]
    g = Duplex(fwd, rev, dimer=Dimer(idx, 3))
    o = ODuplex(fwd, rev, dimer=ODimer(idx, 3))
    assert bool(g) == bool(o) and len(g.dimers) == len(o.dimers) and len(o.dimers) > 0
    for x, y in zip(g.dimers, o.dimers):
        assert x.fwd_matches == y.fwd_matches and abs(x.dG - y.dG) <= 0.01 and x.dG == y.dG
    assert str(g) == str(o)
    # plain construction from two strings (left-aligned)
    g2, o2 = Duplex('GAAGGCAAAGATCGGGAAC', 'GAAGGCTAAGATCGGGAAC', revcomp=True), \
        ODuplex('GAAGGCAAAGATCGGGAAC', 'GAAGGCTAAGATCGGGAAC', revcomp=True)
    assert [d.dG for d in g2.dimers] == [d.dG for d in o2.dimers]
