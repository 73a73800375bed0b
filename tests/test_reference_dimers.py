"""SURVEY §8(f) N3 — self- and cross-dimers of primers: the oracle (CPU) and the CUDA path (-m gpu)
against what the reference's own SecStructures returns (tests/golden/reference_dimers.json.gz, recorded
by tests/golden/make_reference_dimers.py through oracle/refshim.py)."""
import gzip
import json
import os

import numpy as np
import pytest

from oracle import conditions as ocond
from oracle.structures import find_dimers as o_find_dimers

with gzip.open(os.path.join(os.path.dirname(__file__), 'golden', 'reference_dimers.json.gz'), 'rt') as _fh:
    REF = json.load(_fh)
SEQS = REF['seqs']


def mask_of(matches):
    m = 0
    for i in matches:
        m |= 1 << int(i)
    return m


def test_oracle_find_dimers_equals_reference():
    n = 0
    for case in REF['cases']:
        ocond.PCR_P.set(REF['conds'][case['cond']])
        dimers, fronts, n3, min3 = o_find_dimers(SEQS[case['a']], SEQS[case['b']])
        assert [[mask_of(d.fwd_matches), d.offset, d.dG] for d in dimers] == case['dimers']
        assert n3 == case['n_3prim'] and min3 == case['min_3prim_dG']
        if 'fronts' in case:
            assert [[mask_of(d.fwd_matches), d.offset] for d in fronts] == case['fronts']
        n += len(dimers)
    assert n > 200


def _check_gpu(cases, sets):
    n = 0
    for case, ds in zip(cases, sets):
        got = [[mask_of(d.fwd_matches), d.offset] for d in ds.dimers]
        assert got == [x[:2] for x in case['dimers']]
        for d, (_, _, dG) in zip(ds.dimers, case['dimers']):
            assert abs(d.dG - dG) <= 0.01 and abs(d.dG - dG) < 1e-9          # bar / achieved
        assert ds.n_3prim_dimers == case['n_3prim']
        if case['min_3prim_dG'] is None:
            assert ds.min_3prim_dimer_dG is None
        else:
            assert abs(ds.min_3prim_dimer_dG - case['min_3prim_dG']) < 1e-9
        if 'fronts' in case:
            assert [[mask_of(d.fwd_matches), d.offset] for d in ds.fronts] == case['fronts']
        n += len(ds.dimers)
    return n


@pytest.mark.gpu
@pytest.mark.parametrize('cname', sorted(REF['conds']))
def test_gpu_find_dimers_equals_reference(cname):
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SecStructures import find_dimers_batch, find_dimers
    tdf.PCR_P.set(REF['conds'][cname])
    cases = [c for c in REF['cases'] if c['cond'] == cname]
    sets = find_dimers_batch(SEQS, [(c['a'], c['b']) for c in cases])        # one launch for all pairs
    assert _check_gpu(cases, sets) == sum(len(c["dimers"]) for c in cases) > 10
    # the one-pair entry points give the same
    one = [c for c in cases if c['dimers']][:3]
    for c in one:
        ds = find_dimers(SEQS[c['a']], None if c['a'] == c['b'] else SEQS[c['b']])
        _check_gpu([c], [ds])


@pytest.mark.gpu
def test_gpu_find_dimers_equals_oracle_on_random_pairs():
    """Random pairs of unequal length (dangling ends on both sides, long internal loops)."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SecStructures import find_dimers_batch
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.3e-3, DNA=1e-10, DMSO=0.0, PCR_T=40.0)
    tdf.PCR_P.set(cond)
    ocond.PCR_P.set(cond)
    rng = np.random.default_rng(5)
    seqs = []
    for n in list(rng.integers(4, 65, size=40)) + [1, 2, 64, 64]:
        gc = rng.random()
        p = [(1 - gc) / 2, gc / 2, gc / 2, (1 - gc) / 2]
        seqs.append(''.join(rng.choice(list('ACGT'), size=int(n), p=p)))
    pairs = [(int(a), int(b)) for a, b in rng.integers(0, len(seqs), size=(300, 2))] + [(i, i) for i in range(len(seqs))]
    sets = find_dimers_batch(seqs, pairs)
    total = 0
    for (a, b), ds in zip(pairs, sets):
        dimers, fronts, n3, min3 = o_find_dimers(seqs[a], seqs[b])
        assert [(d.fwd_matches, d.offset) for d in ds.dimers] == [(d.fwd_matches, d.offset) for d in dimers]
        assert all(abs(x.dG - y.dG) < 1e-9 for x, y in zip(ds.dimers, dimers))
        assert ds.n_3prim_dimers == n3
        if seqs[a] == seqs[b]:
            assert [(d.fwd_matches, d.offset) for d in ds.fronts] == [(d.fwd_matches, d.offset) for d in fronts]
        total += len(dimers)
    assert total > 100


def test_oracle_find_hairpins_equals_reference():
    from oracle.structures import find_hairpins as o_find_hairpins
    n = 0
    for case in REF['cases']:
        if 'hairpins' not in case:
            continue
        ocond.PCR_P.set(REF['conds'][case['cond']])
        seq = SEQS[case['a']]
        _, fronts, _, _ = o_find_dimers(seq, seq)
        hairpins, n3, _ = o_find_hairpins(fronts, seq)
        assert [[list(h.fwd_matches), list(h.rev_matches), h.dG] for h in hairpins] == case['hairpins']
        assert n3 == case['n_3prim_hairpins']
        n += len(hairpins)
    assert n > 50


@pytest.mark.gpu
@pytest.mark.parametrize('cname', sorted(REF['conds']))
def test_gpu_hairpins_and_secstructures_equal_reference(cname):
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SecStructures import SecStructures, find_dimers_batch, find_hairpins_batch
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    tdf.PCR_P.set(REF['conds'][cname])
    cases = [c for c in REF['cases'] if c['cond'] == cname]
    selfs = [c for c in cases if 'hairpins' in c]
    sets = find_dimers_batch(SEQS, [(c['a'], c['a']) for c in selfs])
    hsets = find_hairpins_batch(sets)                                       # one launch for all seeds
    n = 0
    for c, hs in zip(selfs, hsets):
        assert [[list(h.fwd_matches), list(h.rev_matches)] for h in hs.hairpins] == [x[:2] for x in c['hairpins']]
        for h, (_, _, dG) in zip(hs.hairpins, c['hairpins']):
            assert abs(h.dG - dG) <= 0.01 and abs(h.dG - dG) < 1e-9
        assert hs.n_3prim_hairpins == c['n_3prim_hairpins']
        n += len(hs.hairpins)
    assert n == sum(len(c['hairpins']) for c in selfs) > 10
    # the drop-in class, one by one and batched
    recs = [SeqRecord(Seq(s), id='s%d' % i) for i, s in enumerate(SEQS)]
    some = [c for c in cases if c['dimers'] or c.get('hairpins')][:12] + cases[:4]
    objs = SecStructures.batch(recs, [(c['a'], c['b']) for c in some])
    objs += [SecStructures(recs[c['a']], None if c['a'] == c['b'] else recs[c['b']]) for c in some[:5]]
    for c, o in zip(some + some[:5], objs):
        d = o._seq1_dimers if c['a'] == c['b'] else o._cross_dimers
        assert len(d) == len(c['dimers']) and o._3prim_dimers == c['n_3prim']
        assert sorted(x.dG for x in d.values()) == pytest.approx(sorted(x[2] for x in c['dimers']), abs=1e-9)
        want_min = min((x[2] for x in c['dimers']), default=None)
        assert (o.dimerMin_dG() is None) == (want_min is None)
        if want_min is not None:
            assert abs(o.dimerMin_dG() - want_min) < 1e-9
        if c['a'] == c['b']:
            assert len(o._seq1_hairpins) == len(c['hairpins']) and o._3prim_hairpins == c['n_3prim_hairpins']
            want_h = min((x[2] for x in c['hairpins']), default=None)
            assert (o.hairpinMin_dG() is None) == (want_h is None)
            if want_h is not None:
                assert abs(o.hairpinMin_dG() - want_h) < 1e-9
            assert bool(o) == bool(c['dimers'] or c['hairpins'])
            assert all(h in o for h in o.compose_reactions())
        else:
            assert o.hairpinMin_dG() is None and bool(o) == bool(c['dimers'])


@pytest.mark.gpu
def test_gpu_find_hairpins_equals_oracle_on_random_sequences():
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SecStructures import find_dimers_batch, find_hairpins_batch
    from oracle.structures import find_hairpins as o_find_hairpins
    from oracle.sequtils import reverse_complement
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.3e-3, DNA=1e-10, DMSO=0.0, PCR_T=30.0)
    tdf.PCR_P.set(cond)
    ocond.PCR_P.set(cond)
    rng = np.random.default_rng(11)
    seqs = []
    for n in rng.integers(8, 65, size=60):
        n = int(n)
        stem = ''.join(rng.choice(list('ACGT'), size=int(rng.integers(3, max(4, n // 2 - 2)))))
        loop = ''.join(rng.choice(list('ACGT'), size=int(rng.integers(3, 8))))
        tail = ''.join(rng.choice(list('ACGT'), size=64))
        s = (tail[:int(rng.integers(0, 4))] + stem + loop + reverse_complement(stem) + tail)[:n]
        seqs.append(s)
    sets = find_dimers_batch(seqs, [(i, i) for i in range(len(seqs))])
    hsets = find_hairpins_batch(sets)
    total = 0
    for s, hs in zip(seqs, hsets):
        _, fronts, _, _ = o_find_dimers(s, s)
        hairpins, n3, _ = o_find_hairpins(fronts, s)
        assert [(h.fwd_matches, h.rev_matches) for h in hs.hairpins] == [(h.fwd_matches, h.rev_matches) for h in hairpins]
        assert all(abs(x.dG - y.dG) < 1e-9 for x, y in zip(hs.hairpins, hairpins))
        assert hs.n_3prim_hairpins == n3
        total += len(hairpins)
    assert total > 40


def _facade_from_oracle(cond, a, b):
    """degenprimer_b200.SecStructures.SecStructures filled with the ORACLE's structures (no GPU): what the
    kernels deliver, routed through the same DimerSet / HairpinSet hand-off."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import SecStructures as S
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    from oracle.structures import find_hairpins as o_find_hairpins
    ocond.PCR_P.set(cond)
    tdf.PCR_P.set(cond)
    sa, sb = SEQS[a], SEQS[b]
    dimers, fronts, n3, min3 = o_find_dimers(sa, sb)
    ds = S.DimerSet(sa, sb)
    for d in dimers:
        x = S.Dimer(d.fwd_matches, d.offset)
        x.dG = d.dG
        ds.dimers.append(x)
        if x.fwd_matches[-1] > len(sa) - 4 or x.rev_min < 3:
            ds.n_3prim_dimers += 1
            if ds.min_3prim_dimer_dG is None or ds.min_3prim_dimer_dG > x.dG:
                ds.min_3prim_dimer_dG, ds.min_3prim_dimer = x.dG, x
    assert ds.n_3prim_dimers == n3
    hs = None
    if a == b:
        hs = S.HairpinSet(sa)
        hairpins, hn3, _ = o_find_hairpins(fronts, sa)
        for h in hairpins:
            y = S.Hairpin(h.fwd_matches, h.rev_matches)
            y.dG = h.dG
            hs.hairpins.append(y)
            if y.rev_matches[0] > len(sa) - 4:
                hs.n_3prim_hairpins += 1
                if hs.min_3prim_hairpin_dG is None or hs.min_3prim_hairpin_dG > y.dG:
                    hs.min_3prim_hairpin_dG, hs.min_3prim_hairpin = y.dG, y
        assert hs.n_3prim_hairpins == hn3
    ra, rb = SeqRecord(Seq(sa), id='s%d' % a), SeqRecord(Seq(sb), id='s%d' % b)
    return S.SecStructures(ra, None if a == b else rb, _dimer_set=ds, _hairpin_set=hs)


def test_secstructures_report_text_equals_reference():
    import hashlib
    n_text = n = 0
    for case in REF['cases']:
        obj = _facade_from_oracle(REF['conds'][case['cond']], case['a'], case['b'])
        full, short = str(obj), obj.formatShort()
        if 'text' in case:
            assert [full, short] == case['text']
            n_text += 1
        assert [hashlib.sha1(x.encode()).hexdigest() for x in (full, short)] == case['text_sha1'], \
            (case['cond'], case['a'], case['b'])
        assert obj.formatFull() == full
        n += 1
    assert n_text >= 5 and n == len(REF['cases'])


def _kernel_model_dimers(strs, pairs):
    """What dp_find_dimers writes ([pair][front] = full mask, best mask, best dG), computed with the oracle's
    pieces: the contract of include/degenprimer_b200.h restated on the CPU for the host-side tests."""
    from degenprimer_b200 import _native as nat
    from oracle import thermo
    from oracle.sequtils import reverse_complement
    from oracle.structures import Dimer
    stride = max(len(strs[i]) + len(strs[j]) - 1 for i, j in pairs)
    rec = np.zeros((len(pairs), stride), dtype=nat.FRONT_DIMER_DTYPE)
    for p, (i, j) in enumerate(pairs):
        a, b = strs[i], strs[j]
        rc = reverse_complement(b)
        la, lb = len(a), len(b)
        for front in range(la + lb - 1):
            off = max(0, la - front - 1) - max(0, front - la + 1)
            m = [f for f in range(la) if 0 <= f - off < lb and a[f] == rc[f - off]]
            if not m:
                continue
            d = Dimer(m, off)
            rec[p, front]['full_mask'] = mask_of(m)
            best, best_dG = None, -3
            for x in [d] + d.alternatives:
                g = thermo.dimer_dG_corrected(x, a, b)
                if best_dG > g:
                    best, best_dG = x, g
            if best is not None:
                rec[p, front]['best_mask'] = mask_of(best.fwd_matches)
                rec[p, front]['best_dG'] = best_dG
    return rec


def _kernel_model_hairpins(seqs, seed_seq, seed_mask, seed_off):
    from degenprimer_b200 import _native as nat
    from oracle import thermo
    from oracle.structures import Dimer, Hairpin
    rec = np.zeros(len(seed_seq), dtype=nat.BEST_STRUCTURE_DTYPE)
    for k, (si, mask, off) in enumerate(zip(seed_seq, seed_mask, seed_off)):
        s = seqs[si]
        d = Dimer([i for i in range(64) if mask >> i & 1], off)
        best, best_dG = None, 3
        for x in [d] + d.alternatives:
            h = Hairpin(x.fwd_matches, [len(s) - 1 - (f - off) for f in x.fwd_matches])
            g = thermo.hairpin_dG_corrected(h, s)
            if best_dG > g:
                best, best_dG = x, g
        if best is not None:
            rec[k] = (mask_of(best.fwd_matches), best_dG)
    return rec


@pytest.mark.parametrize('cname', sorted(REF['conds']))
def test_host_halves_of_the_batch_calls_equal_reference(cname):
    """`_dimer_sets_from_records`, `_hairpin_seeds`, `_hairpin_sets_from_records` (everything find_dimers_batch /
    find_hairpins_batch do around the two kernel launches) fed with a CPU model of the kernels' output."""
    from degenprimer_b200 import SecStructures as S
    ocond.PCR_P.set(REF['conds'][cname])
    cases = [c for c in REF['cases'] if c['cond'] == cname]
    pairs = np.array([(c['a'], c['b']) for c in cases], dtype=np.uint32)
    sets = S._dimer_sets_from_records(SEQS, pairs, _kernel_model_dimers(SEQS, pairs))
    assert _check_gpu(cases, sets) == sum(len(c['dimers']) for c in cases)
    selfs = [(c, ds) for c, ds in zip(cases, sets) if 'hairpins' in c]
    seqs = [ds.seq1 for _, ds in selfs]
    seed_seq, seed_mask, seed_off = S._hairpin_seeds([ds for _, ds in selfs])
    hsets = S._hairpin_sets_from_records(seqs, seed_seq, seed_off,
                                         _kernel_model_hairpins(seqs, seed_seq, seed_mask, seed_off))
    for (c, _), hs in zip(selfs, hsets):
        assert [[list(h.fwd_matches), list(h.rev_matches), h.dG] for h in hs.hairpins] == c['hairpins']
        assert hs.n_3prim_hairpins == c['n_3prim_hairpins']
    assert S._hairpin_sets_from_records(['ACGT'], [], [], None)[0].hairpins == []
