"""GPU parity of the C-ABI against the oracle (run on the B200 box: -m gpu)."""
import numpy as np
import pytest

from oracle import search_o1 as o1
from oracle import search_o2 as o2
from oracle import conditions as ocond
from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord
from oracle.sequtils import reverse_complement

from tests.util import random_dna, plant_sites

pytestmark = pytest.mark.gpu

PRIMERS = ['ATATTCTACRACGGCTATCC', 'GAASGCRAAKATYGGGAAC',      # Test_SearchEngine.py:77-78
           'ATARTCTYCGAMGGCTATCC', 'NAAGGYTTAGAKGCGGAAG']        # Test_Product.py:28-29


def split_hits(hits, primer):
    h = hits[hits['primer'] == primer]
    return h[h['strand'] == 0]['pos'].astype(np.int64), h[h['strand'] == 1]['pos'].astype(np.int64)


@pytest.mark.parametrize('algo', [1, 2, 3])
@pytest.mark.parametrize('k', [0, 1, 3, 6])
def test_scan_matches_oracle(ctx, algo, k):
    rng = np.random.default_rng(100 + k)
    t = random_dna(rng, 60000)
    t, _ = plant_sites(rng, t, PRIMERS, min(k, 4), 12)
    tmpl = ctx.template_from_sequences([t])
    assert tmpl.unpack() == t
    assert tmpl.unpack(reverse=True) == reverse_complement(t)
    hits, stats = ctx.scan(tmpl, PRIMERS, k, algo=algo)
    total = 0
    for i, m in enumerate(PRIMERS):
        ref = o2.find(t, m, k)
        f, r = split_hits(hits, i)
        assert np.array_equal(f, ref[4]), (m, 'fwd')
        assert np.array_equal(r, ref[5]), (m, 'rev')
        total += len(f) + len(r)
    assert total == len(hits) and total > 0


@pytest.mark.parametrize('algo', [1, 2, 3])
def test_scan_with_iupac_template_and_fft_oracle(ctx, algo):
    rng = np.random.default_rng(7)
    t = random_dna(rng, 30000, 'ACGTNRYSWKMBDHV', p=[0.23] * 4 + [0.08 / 11] * 11)
    t, _ = plant_sites(rng, t, PRIMERS[:2], 2, 10)
    t = t[:5000] + 'N' * 300 + t[5300:]
    tmpl = ctx.template_from_sequences([t])
    assert tmpl.unpack() == t
    assert tmpl.unpack(reverse=True) == reverse_complement(t)
    for k in (2, 5, 12):
        hits, _ = ctx.scan(tmpl, PRIMERS[:2], k, algo=algo)
        for i, m in enumerate(PRIMERS[:2]):
            ref = o1.find_serial(t, m, k)          # literal FFT restatement
            f, r = split_hits(hits, i)
            assert np.array_equal(f, ref[4]) and np.array_equal(r, ref[5]), (k, m)


def test_multi_record_and_segments(ctx):
    rng = np.random.default_rng(11)
    recs = [random_dna(rng, n) for n in (3000, 64, 20, 4097, 19, 1)]
    recs[0], _ = plant_sites(rng, recs[0], PRIMERS[:1], 1, 4)
    recs[3], _ = plant_sites(rng, recs[3], PRIMERS[:1], 1, 4)
    recs[2] = 'ATATTCTACAACGGCTATCC'                  # exactly one window
    tmpl = ctx.template_from_sequences(recs)
    assert tmpl.n_records == len(recs)
    for r, s in enumerate(recs):
        assert tmpl.unpack(r) == s
    m = PRIMERS[0]
    hits, _ = ctx.scan(tmpl, [m], 1)
    for r, s in enumerate(recs):
        h = hits[hits['record'] == r]
        if len(s) < len(m):
            assert len(h) == 0
            continue
        ref = o2.find(s, m, 1)
        assert np.array_equal(h[h['strand'] == 0]['pos'].astype(np.int64), ref[4])
        assert np.array_equal(h[h['strand'] == 1]['pos'].astype(np.int64), ref[5])
    # sharding by window-start segment reproduces the whole
    cut = tmpl.starts[3] + 2000
    a, _ = ctx.scan(tmpl, [m], 1, 0, cut)
    b, _ = ctx.scan(tmpl, [m], 1, cut, 2 ** 64 - 1)
    both = np.sort(np.concatenate([a, b]), order=['primer', 'strand', 'record', 'pos'])
    assert np.array_equal(both, hits)


def test_errors(ctx):
    with pytest.raises(KeyError):
        ctx.template_from_sequences(['ACGTXACGT'])
    tmpl = ctx.template_from_sequences(['ACGTACGT'])
    with pytest.raises(ValueError):
        ctx.scan(tmpl, ['ACGTACGTACGT'], 1)
    with pytest.raises(ValueError):
        ctx.scan(tmpl, [''], 1)


def _oracle_dimers(t, master, k, cond):
    ocond.PCR_P.set(cond)
    p = OPrimer(OSeqRecord(master, id='p'), 1e-6)
    m = o2.find(t, master, k)
    res = o1.compile_duplexes(m[0], m[1], p, m[2], m[3], m[4], m[5], mp_order=True)
    return p, m, res


@pytest.mark.parametrize('master,k', [('GAASGCRAAKATYGGGAAC', 3), ('ATATTCTACRACGGCTATCC', 2),
                                      ('NAAGGYTTAGAKGCGGAAG', 4)])
def test_duplex_dg_matches_oracle(ctx, master, k):
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    rng = np.random.default_rng(5)
    t = random_dna(rng, 20000, 'ACGTN', p=[0.2495] * 4 + [0.002])
    t, _ = plant_sites(rng, t, [master], k, 20)
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=55.0)
    p, m, ref = _oracle_dimers(t, master, k, cond)
    tdf.PCR_P.set(cond)
    tmpl = ctx.template_from_sequences([t])
    hits, _ = ctx.scan(tmpl, [master], k)
    variants = unambiguous_sequences(master)
    L = len(master)
    dimers, stats = ctx.duplex_dg(tmpl, hits, [L], [len(variants)], [0], ''.join(variants), tdf.conditions())
    # oracle -> flat list of (strand, pos0, variant, mask, dG)
    want = []
    vid = {v: i for i, v in enumerate(variants)}
    for strand, results in enumerate(ref):
        for pos, dups in results:
            for d in dups:
                for dm in d.dimers:
                    mask = sum(1 << i for i in dm.fwd_matches)
                    want.append((strand, pos, vid[d.fwd_seq], mask, dm.dG))
    got = []
    T = len(t)
    for d in dimers:
        h = hits[d['hit']]
        pos = int(h['pos']) + L if h['strand'] == 0 else T + 1 - (int(h['pos']) + L)
        got.append((int(h['strand']), pos, int(d['variant']), int(d['match_mask']), float(d['dG'])))
    assert len(got) == len(want) and len(got) > 0
    key = lambda x: (x[0], x[1], x[2])
    assert sorted(map(key, got)) == sorted(map(key, want))
    # same grouping and the same in-group order (stable sort by dG)
    got_s = sorted(range(len(got)), key=lambda i: (key(got[i]), i))
    want_s = sorted(range(len(want)), key=lambda i: (key(want[i]), i))
    for gi, wi in zip(got_s, want_s):
        assert got[gi][:4] == want[wi][:4]
        assert abs(got[gi][4] - want[wi][4]) <= 1e-9, (got[gi], want[wi])   # bar: 0.01 kcal/mol
    assert max(abs(got[g][4] - want[w][4]) for g, w in zip(got_s, want_s)) == 0.0   # in fact bit-exact


def test_variant_tm_matches_oracle(ctx):
    from degenprimer_b200 import TD_Functions as tdf
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, DMSO=2.0, PCR_T=62.0)
    ocond.PCR_P.set(cond)
    tdf.PCR_P.set(cond)
    p = OPrimer(OSeqRecord('ATARTCTYCGAMGGCNATKCAGGNCTGRGGA', id='p'), 1.5e-6)   # Test_Primer.py:30
    assert p.calculate_Tms()
    from oracle import thermo
    want = np.array([thermo.primer_template_Tm(s, p.concentration) for s in p.str_sequences])
    got = tdf.primer_template_Tms(p.str_sequences, p.concentration)
    assert got.shape == want.shape
    assert np.max(np.abs(got - want)) <= 0.05     # bar from BASELINE.json north_star
    assert np.max(np.abs(got - want)) <= 1e-9


def _two_stage(ctx, tmpl, primers, k, cond, **kw):
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    hits, st = ctx.scan(tmpl, primers, k, **kw)
    var = [unambiguous_sequences(m) for m in primers]
    lens = [len(m) for m in primers]
    offs, blob, off = [], '', 0
    for v, L in zip(var, lens):
        offs.append(off)
        blob += ''.join(v)
        off += len(v) * L
    dimers, dst = ctx.duplex_dg(tmpl, hits, lens, [len(v) for v in var], offs, blob, cond)
    return hits, dimers, var


@pytest.mark.parametrize('k', [1, 3])
def test_find_fused_equals_two_stage(ctx, k):
    """dp_find (one stream submission) == dp_scan + dp_duplex_dg, on a multi-record
    template with N runs and scattered IUPAC codes, primers with different V."""
    from degenprimer_b200 import TD_Functions as tdf
    rng = np.random.default_rng(21 + k)
    recs = [random_dna(rng, n, 'ACGTNRY', p=[0.2485] * 4 + [0.002] * 3) for n in (30000, 500, 12000)]
    recs[0], _ = plant_sites(rng, recs[0], PRIMERS, k, 10)
    recs[2], _ = plant_sites(rng, recs[2], PRIMERS[:2], k, 6)
    recs[0] = recs[0][:9000] + 'N' * 700 + recs[0][9700:]
    cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=52.0)
    tdf.PCR_P.set(cond)
    tmpl = ctx.template_from_sequences(recs)
    hits, dimers, var = _two_stage(ctx, tmpl, PRIMERS, k, tdf.conditions())
    fh, fd, st, dst = ctx.find(tmpl, PRIMERS, k, var, tdf.conditions())
    assert np.array_equal(fh, hits)
    assert np.array_equal(fd, dimers)
    assert len(hits) > 20 and len(dimers) > 20
    assert dst['duplexes'] == sum(len(var[int(h['primer'])]) for h in hits)
    # one record only, through the segment arguments
    b = tmpl.starts[2]
    sh, sd, _, _ = ctx.find(tmpl, PRIMERS, k, var, tdf.conditions(), b, b + 12000)
    sel = hits['record'] == 2
    assert np.array_equal(sh, hits[sel])
    assert len(sd) == int(np.isin(dimers['hit'], np.flatnonzero(sel)).sum())


def test_large_results_and_buffer_growth(ctx):
    """k = L: every window of both strands is a hit (threshold clamp max(1, L-k)), which
    overflows the initial hit buffer (1 Mi records) and the speculative result prefix."""
    from degenprimer_b200 import TD_Functions as tdf
    rng = np.random.default_rng(33)
    T = 600_000
    t = random_dna(rng, T)
    m = 'ACGTACGTACGTACGTAC'
    L = len(m)
    tmpl = ctx.template_from_sequences([t])
    hits, st = ctx.scan(tmpl, [m], L)
    # score >= 1 - 0.5: at least one matching position (or two unknown bases)
    f, r = o2.find(t[:20000], m, L)[4:6]
    n_f = int((hits['strand'] == 0).sum())
    assert n_f > 500_000 and len(hits) > (1 << 20)
    hf = hits[hits['strand'] == 0]['pos'].astype(np.int64)
    assert np.array_equal(hf[hf <= 20000 - L], f)
    assert np.all(np.diff(hf) > 0)
    # fused call on the same data: > 2^20 pairs, dimers mostly filtered out
    tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=30.0))
    fh, fd, _, dst = ctx.find(tmpl, [m], L, [[m]], tdf.conditions())
    assert np.array_equal(fh, hits)
    assert dst['duplexes'] == len(hits)
    sub = hits[:3000]
    d2, _ = ctx.duplex_dg(tmpl, sub, [L], [1], [0], m, tdf.conditions())
    assert np.array_equal(fd[fd['hit'] < 3000], d2)


def test_run_boundary_windows(ctx):
    """Windows that straddle the ends of N runs (1/3 per unknown base) against the
    brute-force oracle: planted sites cut by runs of every short length."""
    rng = np.random.default_rng(77)
    m = PRIMERS[0]
    L = len(m)
    t = list(random_dna(rng, 40000))
    site = 'ATATTCTACAACGGCTATCC'
    at = 100
    for run_len in range(1, 9):
        for cut in (0, 3, L - run_len, L - 1):
            t[at:at + L] = site
            a = at + max(0, min(cut, L - 1))
            t[a:a + run_len] = 'N' * run_len
            at += 150
    # adjacent runs of different letters and a run at the very start / end
    t[at:at + L] = site
    t[at + 5:at + 7] = 'RY'
    t[0:2] = 'NN'
    t[-3:] = 'NNN'
    t = ''.join(t)
    tmpl = ctx.template_from_sequences([t])
    for k in (1, 2, 4):
        ref = o2.find(t, m, k)
        for algo in (1, 2, 3):
            hits, st = ctx.scan(tmpl, [m], k, algo=algo)
            f, r = split_hits(hits, 0)
            assert np.array_equal(f, ref[4]) and np.array_equal(r, ref[5]), (k, algo)
        assert len(ref[4]) > 5


def test_large_duplex_results_device_layout(ctx):
    """More (hit, variant) pairs / dimers than the published prefix holds: the final
    dimer records are laid out by the device (count -> offsets -> gather).  Must equal
    the small path run hit by hit, for dp_find and for dp_duplex_dg."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    rng = np.random.default_rng(91)
    masters = ['ATARTCTYCGAMGGCNATKC', 'GAASGCRAAKATYGGGAAC']      # 64 and 16 variants
    t = random_dna(rng, 120000, 'ACGTN', p=[0.2495] * 4 + [0.002])
    t, _ = plant_sites(rng, t, masters, 3, 60)
    tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=50.0))
    cond = tdf.conditions()
    tmpl = ctx.template_from_sequences([t])
    var = [unambiguous_sequences(m) for m in masters]
    hits, dimers, st, dst = ctx.find(tmpl, masters, 3, var, cond)
    assert dst['duplexes'] > 4096 and len(dimers) > 2048          # the large path was taken
    lens = [len(m) for m in masters]
    offs = [0, len(var[0]) * lens[0]]
    blob = ''.join(var[0]) + ''.join(var[1])
    d2, _ = ctx.duplex_dg(tmpl, hits, lens, [len(v) for v in var], offs, blob, cond)
    assert np.array_equal(d2, dimers)
    # reference: small calls of 2 hits each (published-prefix path), indices shifted
    parts = []
    for a in range(0, len(hits), 2):
        d, s = ctx.duplex_dg(tmpl, hits[a:a + 2], lens, [len(v) for v in var], offs, blob, cond)
        assert s['duplexes'] <= 4096 and len(d) <= 2048
        d = d.copy()
        d['hit'] += a
        parts.append(d)
    want = np.concatenate(parts)
    assert np.array_equal(want, dimers)
    assert np.all(np.diff(dimers['hit'].astype(np.int64)) >= 0)


def test_jit_kernels_equal_generic(ctx):
    """Run-time specialised scan kernels (NVRTC): same hits as the generic bit-slice kernel
    and the popcount kernel for plain, degenerate and many-primer sets; prepare() makes
    DP_ALGO_AUTO pick them up."""
    from degenprimer_b200 import _native as nat
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    from degenprimer_b200.synthetic import multiplex_primers
    rng = np.random.default_rng(123)
    sets = [(PRIMERS, 1), (['AGRGTTYGATYMTGGYTYWG', 'KACGGYTAMCTTRTTAYGASTW'], 3),
            (multiplex_primers(), 3), (['ATATTCTACAACGGCTATCCNNNNNNAGGTCA', 'GATTACAGATTACAGATTACA'], 4)]
    for masters, k in sets:
        t = random_dna(rng, 300000, 'ACGTN', p=[0.2495] * 4 + [0.002])
        t, _ = plant_sites(rng, t, masters, min(k, 3), 6)
        tmpl = ctx.template_from_sequences([t])
        a, sa = ctx.scan(tmpl, masters, k, algo=nat.DP_ALGO_JIT)
        b, sb = ctx.scan(tmpl, masters, k, algo=nat.DP_ALGO_BITSLICE)
        c, sc = ctx.scan(tmpl, masters, k, algo=nat.DP_ALGO_POPC)
        assert sa['algo'] == nat.DP_ALGO_JIT and sb['algo'] == nat.DP_ALGO_BITSLICE
        assert np.array_equal(a, b) and np.array_equal(a, c) and len(a) > 0
        d, sd = ctx.scan(tmpl, masters, k)                       # compiled above: AUTO uses it, whatever the set size
        assert sd['algo'] == nat.DP_ALGO_JIT and np.array_equal(a, d)
    # a set nobody prepared runs on the generic kernel under AUTO; prepare() switches it
    fresh = ['TTGACCGATAGGCTAACGTA']
    t = random_dna(rng, 50000)
    tmpl = ctx.template_from_sequences([t])
    _, s0 = ctx.scan(tmpl, fresh, 2)
    assert s0['algo'] == nat.DP_ALGO_BITSLICE
    q = ctx.prepare(fresh, 2, [unambiguous_sequences(m) for m in fresh], jit=True)
    assert q.jit, getattr(q, 'jit_error', '')
    _, s1 = ctx.scan(tmpl, fresh, 2)
    assert s1['algo'] == nat.DP_ALGO_JIT


def test_find_again_under_new_conditions(ctx):
    """dp_find_again: the duplex stage of the last find() under other PCR conditions, hits
    still on the device, equals a fresh find() under those conditions — for a result that
    fits the published prefix and for one laid out by the device — and refuses to run once
    another call has used the context."""
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import _native as nat
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    rng = np.random.default_rng(404)
    first = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=52.0)
    second = dict(Na=120e-3, Mg=1.5e-3, dNTP=0.2e-3, DNA=1e-9, PCR_T=64.0)
    for masters, k, n_sites in ((PRIMERS[:2], 2, 12), (['ATARTCTYCGAMGGCNATKC', 'GAASGCRAAKATYGGGAAC'], 3, 60)):
        t = random_dna(rng, 150000, 'ACGTN', p=[0.2495] * 4 + [0.002])
        t, _ = plant_sites(rng, t, masters, min(k, 3), n_sites)
        tmpl = ctx.template_from_sequences([t])
        var = [unambiguous_sequences(m) for m in masters]
        q = ctx.prepare(masters, k, var)
        tdf.PCR_P.set(first)
        c1 = tdf.conditions()
        hits, d1, _, _ = ctx.find(tmpl, q, cond=c1)
        tdf.PCR_P.set(second)
        c2 = tdf.conditions()
        again, dst = ctx.find_again(c2)
        hits2, d2, _, dst2 = ctx.find(tmpl, q, cond=c2)
        assert np.array_equal(hits, hits2) and len(hits) > 10
        assert np.array_equal(again, d2) and dst['duplexes'] == dst2['duplexes']
        assert not np.array_equal(d1, d2)                         # the conditions matter
        back, _ = ctx.find_again(c1)                              # and back: same records as the first call
        assert np.array_equal(back, d1)
        ctx.scan(tmpl, masters, k)                                # another call takes the hit workspace
        with pytest.raises(nat.NativeError) as err:
            ctx.find_again(c2)
        assert err.value.code == nat.DP_ERR_STATE
    tdf.PCR_P.set(first)
