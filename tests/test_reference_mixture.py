"""SURVEY §8(f) N1 — product pairing: degenprimer_b200.PCR_Mixture.MixtureFactory against what the
reference's own MixtureFactory.create_PCR_mixture returned (tests/golden/reference_mixture.json.gz,
recorded by tests/golden/make_reference_mixture.py through oracle/refshim.py).  Host code: the CPU test
feeds it the oracle's annealings, the -m gpu test the CUDA path's."""
import gzip
import json
import os
import pickle
import threading

import pytest

from oracle import conditions as ocond
from oracle import search_o1 as o1
from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord

from degenprimer_b200.PCR_Mixture import MixtureFactory, load_mixture
from degenprimer_b200.Product import Product, Region
from degenprimer_b200.WorkCounter import WorkCounter

with gzip.open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_mixture.json.gz'), 'rt') as _fh:
    REF = json.load(_fh)


def region(r):
    return [r.name, r.start, r.end, bool(r.forward)]


def dump(mix):
    if mix is None:
        return None
    return {'reaction_id': mix.reaction_id,
            'products': sorted([p.start, p.end, p.fwd_ids, p.rev_ids, region(p.fwd_template), region(p.rev_template),
                                len(p)] for p in mix.products.values()),
            'fwd_templates': [region(r) for r in mix.templates[1]],
            'rev_templates': [region(r) for r in mix.templates[0]],
            'annealings': [[region(t), [d.name for d in dups]] for dups, t in mix.annealings]}


@pytest.mark.parametrize('idx', range(len(REF['cases'])))
def test_mixture_factory_equals_reference(idx):
    case = REF['cases'][idx]
    ocond.PCR_P.set(REF['conds'][case['cond']])
    primers = [OPrimer(OSeqRecord(m, id=pid), 1e-6) for pid, m in case['primers']]
    fwd, rev = [], []
    for p, (_, m) in zip(primers, case['primers']):
        r = o1.find_serial(case['template'], m, case['k'])
        ann = o1.compile_duplexes(r[0], r[1], p, r[2], r[3], r[4], r[5], mp_order=False)
        if ann is None:
            continue
        fwd.extend(ann[0] or [])
        rev.extend(ann[1] or [])
    assert [len(fwd), len(rev)] == case['n_sites']
    mf = MixtureFactory(threading.Event(), primers, case['min_amplicon'], case['max_amplicon'],
                        with_exonuclease=case['with_exonuclease'],
                        include_side_annealings=case['include_side_annealings'])
    c = WorkCounter()
    mix = mf.create_PCR_mixture(c, 'T synthetic', fwd, rev)
    assert json.loads(json.dumps(dump(mix))) == case['mixture']
    if mix is not None:
        again = load_mixture(mix.save())                       # survives the temp-file hand-off
        assert json.loads(json.dumps(dump(again))) == case['mixture']
        p = next(iter(mix.products.values()))
        text = p.pretty_print()
        assert 'forward annealing site' in text and 'concentration:    0.0  M' in text and str(p)


def test_region_and_product_semantics():
    a, b = Region('t', 10, 20, True), Region('t', 15, 30, True)
    assert a.overlaps(b) and len(a) == 11 and a != b and a == Region('t', 10, 20, True)
    a += b
    assert (a.start, a.end) == (10, 30)
    a += Region('other', 1, 99, True)                          # another sequence: ignored
    assert (a.start, a.end) == (10, 30)
    assert Region('x', 1, 5, False).overlaps(Region('y', 2, 3, True))      # the reference's precedence slip: end inside is enough
    with pytest.raises(ValueError):
        Region('t', 0, 5, True)
    with pytest.raises(ValueError):
        Region('t', 6, 5, True)

    class Dup(object):
        def __init__(self, name, n):
            self.name, self.fwd_len = name, n
    d1, d2, d3 = Dup('P_1', 20), Dup('P_2', 25), Dup('R_1', 18)
    p = Product('t', 101, 300, [d1], [d3])
    assert region(p.fwd_template) == ['t', 81, 100, True] and region(p.rev_template) == ['t', 301, 318, False]
    q = Product('t', 101, 300, [d2], [d3])
    p += q
    assert p.fwd_ids == ['P_1', 'P_2'] and region(p.fwd_template) == ['t', 76, 100, True]
    wider = Product('t', 90, 300, [d2], [d3])
    p += wider
    assert p.start == 90 and p.fwd_ids == ['P_2'] and hash(p) == hash(Product('t', 90, 300))
    assert region(Product('t', 1, 50, [d1]).fwd_template) == ['t', 1, 1, True]      # clamped at the sequence start
    p2 = pickle.loads(pickle.dumps(Product('t', 90, 300)))
    assert p2 == p and region(p2.rev_template) == ['t', 301, 301, False]


@pytest.mark.gpu
@pytest.mark.parametrize('idx', range(len(REF['cases'])))
def test_gpu_annealings_to_mixture_equal_reference(idx):
    """The CUDA path's annealings through the same factory: scan + dG + pairing == the reference end to end."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    case = REF['cases'][idx]
    tdf.PCR_P.set(REF['conds'][case['cond']])
    primers = []
    for pid, m in case['primers']:
        p = Primer(SeqRecord(Seq(m), id=pid), 1e-6)
        p.generate_components()
        primers.append(p)
    se = SearchEngine(threading.Event())
    rec = SeqRecord(Seq(case['template']), id='T', description='T synthetic')
    fwd, rev = [], []
    for p in primers:
        m = se.find_matches(WorkCounter(), rec, p, case['k'])
        ann = se.compile_duplexes(WorkCounter(), *m)
        if ann is None:
            continue
        fwd.extend(ann[0] or [])
        rev.extend(ann[1] or [])
    assert [len(fwd), len(rev)] == case['n_sites']
    mf = MixtureFactory(threading.Event(), primers, case['min_amplicon'], case['max_amplicon'],
                        with_exonuclease=case['with_exonuclease'],
                        include_side_annealings=case['include_side_annealings'])
    mix = mf.create_PCR_mixture(WorkCounter(), 'T synthetic', fwd, rev)
    assert json.loads(json.dumps(dump(mix))) == case['mixture']


@pytest.mark.parametrize('idx', range(len(REF['cases'])))
def test_facade_site_objects_through_the_factory(idx):
    """The facade's lazily materialised result objects (DuplexTable / SiteDuplexes / Duplex._from_device: what
    the CUDA path hands out) carry everything the pairing reads — built here from the oracle's numbers, no GPU."""
    import numpy as np
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import _native as nat
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SecStructures import DuplexTable, SiteDuplexes
    from degenprimer_b200.SeqUtils import SeqRecord, Seq, reverse_complement
    case = REF['cases'][idx]
    ocond.PCR_P.set(REF['conds'][case['cond']])
    tdf.PCR_P.set(REF['conds'][case['cond']])
    primers, fwd, rev = [], [], []
    for pid, m in case['primers']:
        gp = Primer(SeqRecord(Seq(m), id=pid), 1e-6)
        gp.generate_components()
        primers.append(gp)
        op = OPrimer(OSeqRecord(m, id=pid), 1e-6)
        r = o1.find_serial(case['template'], m, case['k'])
        ann = o1.compile_duplexes(r[0], r[1], op, r[2], r[3], r[4], r[5], mp_order=False)
        if ann is None:
            continue
        names = [x.id for x in gp.seq_records]
        for strand_list, sink in ((ann[0] or [], fwd), (ann[1] or [], rev)):
            rows, windows, spans = [], [], []
            for pos, dups in strand_list:
                h = len(windows)
                windows.append(reverse_complement(dups[0].rev_seq))
                lo = len(rows)
                for d in dups:
                    v = names.index(d.name)
                    for dm in d.dimers:
                        mask = 0
                        for i in dm.fwd_matches:
                            mask |= 1 << i
                        rows.append((h, v, mask, dm.dG))
                spans.append((pos, h, lo, len(rows)))
            table = DuplexTable(np.array(rows, dtype=nat.DIMER_DTYPE), windows, gp.str_sequences, names,
                                tdf.PCR_P.PCR_T)
            sink.extend((pos, SiteDuplexes(table, h, lo, hi)) for pos, h, lo, hi in spans)
    mf = MixtureFactory(threading.Event(), primers, case['min_amplicon'], case['max_amplicon'],
                        with_exonuclease=case['with_exonuclease'],
                        include_side_annealings=case['include_side_annealings'])
    mix = mf.create_PCR_mixture(WorkCounter(), 'T synthetic', fwd, rev)
    assert json.loads(json.dumps(dump(mix))) == case['mixture']
    if mix is not None:
        assert json.loads(json.dumps(dump(load_mixture(mix.save())))) == case['mixture']


class _OracleSearcher(object):
    """Stands in for the CUDA SearchEngine in the composition test below: same four methods, oracle inside."""

    def __init__(self, oprimers):
        self.oprimers = oprimers                      # master string -> oracle Primer

    def find_matches(self, counter, template, primer, mismatches):
        r = o1.find_matches(str(template.seq), str(primer), mismatches, workers=0)
        return (r[0], r[1], primer, r[2], r[3], r[4], r[5])

    def _compile(self, matches, mp):
        f, rv, primer, t_len, p_len, fm, rm = matches
        return o1.compile_duplexes(f, rv, self.oprimers[str(primer)], t_len, p_len, fm, rm, mp_order=mp)

    def compile_duplexes(self, counter, *matches):
        return self._compile(matches, False)

    def compile_duplexes_mp(self, counter, *matches):
        return self._compile(matches, True)

    def find(self, counter, template, primer, mismatches):
        from degenprimer_b200.tmpStorage import to_shelf
        res = self._compile(self.find_matches(counter, template, primer, mismatches), True)
        return to_shelf(res) if res else None

    def find_all(self, counter, templates, primers, mismatches):
        return {t.id: [self._compile(self.find_matches(counter, t, p, mismatches), True) for p in primers]
                for t in templates}


@pytest.mark.parametrize('idx', [0, 2, 5])
def test_products_finder_composition(idx):
    """PCR_ProductsFinder.find / find_matches + matches_to_mixture / batch_find give the reference's mixture
    when the searcher gives the reference's annealings (host logic only; the searcher is the oracle here)."""
    from degenprimer_b200.PCR_ProductsFinder import PCR_ProductsFinder
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    case = REF['cases'][idx]
    ocond.PCR_P.set(REF['conds'][case['cond']])
    primers = []
    for pid, m in case['primers']:
        p = Primer(SeqRecord(Seq(m), id=pid), 1e-6)
        p.generate_components()
        primers.append(p)
    ppf = PCR_ProductsFinder(threading.Event(), primers, case['min_amplicon'], case['max_amplicon'],
                             with_exonuclease=case['with_exonuclease'],
                             include_side_annealings=case['include_side_annealings'])
    ppf._searcher = _OracleSearcher({m: OPrimer(OSeqRecord(m, id=pid), 1e-6) for pid, m in case['primers']})
    rec = SeqRecord(Seq(case['template']), id='T', description='T synthetic')
    want = case['mixture']
    # the serial pooling order of the fixtures == batch_find's (compile_duplexes per primer)
    res = ppf.batch_find(WorkCounter(), [rec], case['k'])
    if want is None:
        assert res is None and ppf.find(WorkCounter(), rec, case['k']) is None
        return
    assert list(res) == ['synthetic']                          # pretty_rec_name: description without the id
    got = dump(load_mixture(res['synthetic']))
    got['reaction_id'] = want['reaction_id']                   # the fixtures were recorded under 'T synthetic'
    for key in ('products', 'fwd_templates', 'rev_templates', 'annealings'):
        got[key] = json.loads(json.dumps(got[key]).replace('"synthetic"', '"T synthetic"'))
    assert got == want
    # find() and find_matches() + matches_to_mixture() pool position-sorted lists: same products
    c = WorkCounter()
    res2 = ppf.find(c, rec, case['k'])
    matches = ppf.find_matches(WorkCounter(), rec, case['k'])
    assert len(matches) == len(primers) and matches[0][2] is primers[0]
    path3 = ppf.matches_to_mixture(WorkCounter(), 'synthetic', matches)
    for path in (res2['synthetic'], path3):
        mix = load_mixture(path)
        prods = json.loads(json.dumps(dump(mix)['products']).replace('"synthetic"', '"T synthetic"'))
        assert prods == want['products']
    ev = threading.Event()
    ev.set()
    ppf._abort_event = ev
    assert ppf.find(WorkCounter(), rec, case['k']) is None and ppf.find_matches(WorkCounter(), rec, case['k']) is None


def test_product_text_equals_reference():
    from tests.golden.make_reference_mixture import StubDuplex
    want = REF['texts']
    a = Product('t seq', 101, 300, [StubDuplex('P_1', 20), StubDuplex('P_2', 25)], [StubDuplex('R_1', 18)])
    a.quantity, a.cycles = 2.5e-7, 12
    assert str(a) == want['str'] and a.pretty_print() == want['pretty']
    assert a.pretty_print(False, False) == want['pretty_no_name_no_3']
    assert str(Region('x', 3, 9, False)) == want['region'] and a.fwd_ids == want['fwd_ids']
    assert [a.fwd_template.start, a.fwd_template.end] == want['fwd_template']


def test_reference_product_smoke_test():
    """Tests/DegenPrimer_Tests/Test_Product.py:27-39 — the one assertion the reference's tests make on this
    side of the path: a Product of two duplexes survives save/load with the same text.  Here also: that text
    equals the reference's (recorded), with the oracle's Duplex (degenerate primers: averaged NN parameters)."""
    from oracle.sequtils import reverse_complement
    from oracle.structures import Duplex
    want = REF['test_product']
    ocond.PCR_P.set(REF['conds'][want['cond']])
    d1 = Duplex(want['fwd'], reverse_complement(want['fwd']))
    d2 = Duplex(want['rev'], reverse_complement(want['rev']))
    assert [d1.dG, d2.dG] == want['dG']
    p = Product('test_template', 50, 150, (d1,), (d2,))
    p1 = pickle.loads(pickle.dumps(p))
    assert str(p) == str(p1) == want['str'] and p.pretty_print() == p1.pretty_print() == want['pretty']
