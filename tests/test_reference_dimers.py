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
