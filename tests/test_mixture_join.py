"""SURVEY §8(f) N1 as a device join: dp_pair_products (the pairing loop of MixtureFactory.create_PCR_mixture with
its rev_start quirk), the array-only good / bad 3' split over SitePools and the lazily materialised mixture.
CPU: the array path fed with the oracle's annealings (the pairing kernel stood in by the literal Python loop)
gives the reference's recorded mixtures.  -m gpu: the kernel against that loop on random site systems, and the
CUDA path end to end through PCR_ProductsFinder.find on the recorded systems."""
import json
import threading

import numpy as np
import pytest

from degenprimer_b200 import _native as nat
from degenprimer_b200.PCR_Mixture import LazyMixture, MixtureFactory, load_mixture
from degenprimer_b200.SecStructures import DuplexTable, SiteList, SitePool
from degenprimer_b200.WorkCounter import WorkCounter
from oracle import conditions as ocond
from oracle import search_o1 as o1
from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord
from oracle.sequtils import reverse_complement
from tests.test_reference_mixture import REF, dump


def loop_pairs(fwd_pos, rev_pos, lo, hi):
    """DegenPrimer/PCR_Mixture.py:216-236, literally: (pairs, slice start per forward site)."""
    pairs, starts, rev_start = [], [], 0
    for f, fp in enumerate(fwd_pos):
        start = fp + 1
        starts.append(rev_start)
        base = rev_start
        for rel, rp in enumerate(rev_pos[base:]):
            end = rp - 1
            if start >= end:
                rev_start = rel
                continue
            if not lo <= end - start + 1 <= hi:
                break
            pairs.append((f, base + rel))
    return pairs, starts


class _LoopCtx(object):
    """Stands in for the native context in the CPU test: same call, the loop above inside."""

    def pair_products(self, fwd_pos, rev_pos, lo, hi):
        pairs, starts = loop_pairs([int(x) for x in fwd_pos], [int(x) for x in rev_pos], lo, hi)
        arr = np.array(pairs, dtype=nat.PAIR_DTYPE) if pairs else np.zeros(0, dtype=nat.PAIR_DTYPE)
        return arr, np.array(starts, dtype=np.uint32)


def site_list_from_oracle(annealings, primer, T):
    """[(pos, (oracle Duplex, ...)), ...] of one primer and strand -> SiteList over a DuplexTable."""
    variants = [str(r.seq) for r in primer.seq_records]
    names = [r.id for r in primer.seq_records]
    rows, windows, pos, lo, hi = [], [], [], [], []
    for h, (p, dups) in enumerate(annealings):
        lo.append(len(rows))
        for d in dups:
            v = variants.index(d._fwd_sequence)
            for x in d.dimers:
                rows.append((h, v, sum(1 << i for i in x.fwd_matches), x.dG))
        hi.append(len(rows))
        windows.append(reverse_complement(dups[0]._rev_sequence))
        pos.append(p)
    table = DuplexTable(np.array(rows, dtype=nat.DIMER_DTYPE), windows, variants, names, T)
    return SiteList(table, np.arange(len(pos)), np.array(pos, dtype=np.int64), np.array(lo), np.array(hi))


@pytest.mark.parametrize('idx', range(len(REF['cases'])))
def test_array_path_gives_the_reference_mixture(idx, monkeypatch):
    from degenprimer_b200 import TD_Functions as tdf
    case = REF['cases'][idx]
    ocond.PCR_P.set(REF['conds'][case['cond']])
    tdf.PCR_P.set(REF['conds'][case['cond']])
    monkeypatch.setattr(tdf, '_ctx', lambda: _LoopCtx())
    primers = [OPrimer(OSeqRecord(m, id=pid), 1e-6) for pid, m in case['primers']]
    fwd, rev = SitePool(), SitePool()
    for p, (_, m) in zip(primers, case['primers']):
        r = o1.find_serial(case['template'], m, case['k'])
        ann = o1.compile_duplexes(r[0], r[1], p, r[2], r[3], r[4], r[5], mp_order=False)
        if ann is None:
            continue
        for pool, lst in ((fwd, ann[0]), (rev, ann[1])):
            if lst:
                pool.extend(site_list_from_oracle(lst, p, tdf.PCR_P.PCR_T))
    assert [len(fwd), len(rev)] == case['n_sites']
    mf = MixtureFactory(threading.Event(), primers, case['min_amplicon'], case['max_amplicon'],
                        with_exonuclease=case['with_exonuclease'],
                        include_side_annealings=case['include_side_annealings'])
    mix = mf.create_PCR_mixture(WorkCounter(), 'T synthetic', fwd, rev)
    if case['mixture'] is None:
        assert mix is None
        return
    assert isinstance(mix, LazyMixture) and mix._real is None and mix.n_products >= len(case['mixture']['products'])
    starts, ends = mix.product_bounds()
    assert sorted(set(zip(starts.tolist(), ends.tolist()))) == sorted({(p[0], p[1]) for p in case['mixture']['products']})
    again = load_mixture(mix.save())                          # pickled as arrays, materialised after the hand-off
    assert isinstance(again, LazyMixture) and again._real is None
    assert json.loads(json.dumps(dump(again))) == case['mixture']
    assert json.loads(json.dumps(dump(mix))) == case['mixture']


def test_loop_stand_in_has_the_quirks():
    """The literal loop the kernel is compared with: rev_start is an index into the slice, a too short first
    candidate ends the scan."""
    pairs, starts = loop_pairs([10, 20, 300], [5, 8, 15, 100, 200, 400], 50, 500)
    assert starts == [0, 1, 1] and pairs == [(1, 3), (1, 4), (1, 5), (2, 5)]       # site 0: 15 is too close -> break
    assert loop_pairs([10], [30, 100], 50, 500)[0] == []          # 30 is too close: break, 100 is never looked at


@pytest.mark.gpu
def test_pair_products_kernel_equals_the_loop(ctx):
    rng = np.random.default_rng(7)
    total = 0
    for trial in range(300):
        n_f, n_r = int(rng.integers(0, 60)), int(rng.integers(0, 60))
        span = int(rng.choice([200, 2000, 20000]))
        fwd = np.sort(rng.integers(1, span, size=n_f))
        rev = np.sort(rng.integers(1, span, size=n_r))
        lo = int(rng.integers(1, 200))
        hi = lo + int(rng.integers(0, 3000))
        pairs, starts = ctx.pair_products(fwd, rev, lo, hi)
        want, wstarts = loop_pairs(fwd.tolist(), rev.tolist(), lo, hi)
        assert list(zip(pairs['fwd'].tolist(), pairs['rev'].tolist())) == want, trial
        assert starts.tolist() == wstarts
        total += len(want)
    assert total > 5000
    # a large system: 50,000 forward and reverse sites
    fwd = np.sort(rng.integers(1, 5_000_000, size=50_000))
    rev = np.sort(rng.integers(1, 5_000_000, size=50_000))
    pairs, starts = ctx.pair_products(fwd, rev, 100, 3000)
    want, wstarts = loop_pairs(fwd.tolist(), rev.tolist(), 100, 3000)
    assert len(want) > 100_000 and list(zip(pairs['fwd'].tolist(), pairs['rev'].tolist())) == want
    assert starts.tolist() == wstarts
    with pytest.raises(nat.NativeError, match='not sorted'):
        ctx.pair_products([5, 3], [10], 1, 100)


@pytest.mark.gpu
@pytest.mark.parametrize('idx', range(len(REF['cases'])))
def test_products_finder_find_on_the_device_join(idx):
    """PCR_ProductsFinder.find on the CUDA path: one dp_find for all primers, SitePools, dp_pair_products,
    lazily materialised mixture == the reference's products."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.PCR_ProductsFinder import PCR_ProductsFinder
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqUtils import Seq, SeqRecord
    case = REF['cases'][idx]
    tdf.PCR_P.set(REF['conds'][case['cond']])
    primers = []
    for pid, m in case['primers']:
        p = Primer(SeqRecord(Seq(m), id=pid), 1e-6)
        p.generate_components()
        primers.append(p)
    ppf = PCR_ProductsFinder(threading.Event(), primers, case['min_amplicon'], case['max_amplicon'],
                             with_exonuclease=case['with_exonuclease'],
                             include_side_annealings=case['include_side_annealings'])
    rec = SeqRecord(Seq(case['template']), id='T', description='T synthetic')
    res = ppf.find(WorkCounter(), rec, case['k'])
    if case['mixture'] is None:
        assert res is None
        return
    mix = load_mixture(res['synthetic'])
    assert isinstance(mix, LazyMixture)
    got = dump(mix)
    prods = json.loads(json.dumps(got['products']).replace('"synthetic"', '"T synthetic"'))
    assert prods == case['mixture']['products']              # find() pools position-sorted lists: same products
    assert len(got['annealings']) == len(case['mixture']['annealings'])
