"""Round-2 GPU tests (-m gpu): the other BASELINE configs at FULL size against the C brute-force oracle
(stretches + every planted site) and the Python oracle's duplexes, the in-memory packed image, the
cache identity, `SearchEngine.find_all`, and the genome sharding on one GPU."""
import os
import threading

import numpy as np
import pytest

from oracle import conditions as ocond
from oracle import o2c
from oracle import search_o1 as o1
from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord

pytestmark = pytest.mark.gpu

COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0)


def _variants(primers):
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    return [unambiguous_sequences(p) for p in primers]


def _conditions():
    from degenprimer_b200 import TD_Functions as tdf
    ocond.PCR_P.set(COND)
    tdf.PCR_P.set(COND)
    return tdf.conditions()


def _check_stretch(hits, records, lengths, primers, k, rec, a, b):
    """GPU hits whose window lies in bases [a, b) of record `rec` == the C oracle's on those bytes."""
    from bench import gpu_sites_in
    text = records[rec][a:b].tobytes()
    got = set(gpu_sites_in(hits, None, rec, a, b, [len(p) for p in primers], lengths[rec]))
    want = set()
    for pi, m in enumerate(primers):
        f, r = o2c.find(text, m, k)
        want.update((pi, 0, int(x)) for x in f)
        want.update((pi, 1, int(x)) for x in r)
    assert got == want, (rec, a, b, sorted(got ^ want)[:5])
    return len(want)


def _check_planted(hits, wl, primers, k):
    """Every planted site that still scores (a later plant or an N run may have overwritten it) is a hit."""
    found = {(int(h['record']), int(h['primer']), int(h['strand']), int(h['pos'])) for h in hits}
    n = 0
    for rec, planted in enumerate(wl['planted']):
        arr = wl['records'][rec]
        if arr is None:
            continue
        T = len(arr)
        for pi, strand, at in planted:
            L = len(primers[pi])
            f, r = o2c.find(arr[at:at + L].tobytes(), primers[pi], k)
            if len(f if strand == 0 else r):
                pos = at if strand == 0 else T - L - at
                assert (rec, pi, strand, pos) in found, (rec, pi, strand, at)
                n += 1
    return n


def _check_dimers(hits, dimers, wl, primers, var_lists, rec, a, b):
    """Dimers of the GPU hits inside a stretch == the oracle's Duplex objects on the same windows."""
    from bench import gpu_sites_in
    arr = wl['records'][rec]
    got = gpu_sites_in(hits, dimers, rec, a, b, [len(p) for p in primers], len(arr))
    text = arr[a:b].tobytes().decode()
    from oracle.sequtils import reverse_complement
    rev = reverse_complement(text)
    n = 0
    for (pi, strand, pos), rows in got.items():
        primer = OPrimer(OSeqRecord(primers[pi], id='p'), 1.5e-6)
        with ocond.AcquireParameters():
            _, dups = o1.compile_duplexes_for_position(pos, rev if strand else text, primer, len(text),
                                                       len(primers[pi]), bool(strand))
        want = []
        for d in dups:
            v = var_lists[pi].index(d._fwd_sequence)
            for x in d.dimers:
                want.append((v, sum(1 << i for i in x.fwd_matches), x.dG))
        assert [(v, m) for v, m, _ in rows] == [(v, m) for v, m, _ in want], (pi, strand, pos)
        for (_, _, g1), (_, _, g2) in zip(rows, want):
            assert abs(g1 - g2) <= 0.01            # bar of BASELINE.json; the kernels are in fact bit-exact
            n += 1
    return n


@pytest.fixture(scope='module')
def genome_3g(ctx):
    """The 3.1 Gbp, 24-record genome of config 4 and the north-star pair, resident once."""
    from degenprimer_b200.synthetic import make_workload
    wl = make_workload('config4_20plex_3Gbp')
    tmpl = ctx.template_from_sequences([(r.ctypes.data, len(r)) for r in wl['records']])
    yield wl, tmpl
    tmpl.close()


def test_config4_full_size(ctx, genome_3g):
    """20-plex, k = 3, 3.1 Gbp: generic and run-time specialised kernels agree, two 20 Mbp stretches ==
    C oracle for all 20 primers, every planted site is found, dimers of a stretch == oracle Duplexes."""
    wl, tmpl = genome_3g
    assert tmpl.total_bases == 3_099_999_988 and tmpl.n_records == 24
    assert tmpl.n_exception_runs == 2 * sum(1 for n in wl['lengths'] if n > 8_000_010)      # two runs of N per record
    primers, k = wl['primers'], wl['k']
    var = _variants(primers)
    cond = _conditions()
    hits, dimers, st, dst = ctx.find(tmpl, ctx.prepare(primers, k, var), cond=cond, algo=2)
    assert st['algo'] == 2 and len(hits) > 1500
    hj, dj, stj, _ = ctx.find(tmpl, ctx.prepare(primers, k, var, jit=True), cond=cond, algo=3)
    assert stj['algo'] == 3
    assert np.array_equal(hits, hj) and np.array_equal(dimers, dj)
    assert _check_stretch(hits, wl['records'], wl['lengths'], primers, k, 2, 40_000_000, 60_000_000) > 0
    assert _check_stretch(hits, wl['records'], wl['lengths'], primers, k, 23, 0, 20_000_000) > 0
    assert _check_planted(hits, wl, primers, k) > 1500
    a = wl['planted'][5][0][2] // 32 * 32
    assert _check_dimers(hits, dimers, wl, primers, var, 5, max(0, a - 500_000), a + 500_000) > 0


def test_northstar_full_size(ctx, genome_3g):
    """256-fold degenerate pair, k = 3, on the same genome: stretch == C oracle, planted sites found,
    every dimer of every (hit, variant) of a stretch == the oracle's Duplex objects."""
    from degenprimer_b200.synthetic import P27F_D128, P1492R_D128, make_record
    wl, tmpl = genome_3g
    primers, k = [P27F_D128, P1492R_D128], 3
    var = _variants(primers)
    assert sum(map(len, var)) == 256
    cond = _conditions()
    hits, dimers, st, dst = ctx.find(tmpl, ctx.prepare(primers, k, var, jit=True), cond=cond)
    assert st['algo'] == 3 and dst['duplexes'] == 128 * len(hits) and len(hits) > 5000
    h2, d2, st2, _ = ctx.find(tmpl, ctx.prepare(primers, k, var), cond=cond, algo=2)
    assert np.array_equal(hits, h2) and np.array_equal(dimers, d2)
    assert _check_stretch(hits, wl['records'], wl['lengths'], primers, k, 14, 100_000_000, 130_000_000) > 0
    # planted sites of the pair (make_record returns the sites of the workload's own primers)
    ns = {'records': wl['records'], 'planted': [make_record('northstar_256fold_3Gbp', i, n)[1] if i in (0, 7, 23) else []
                                                for i, n in enumerate(wl['lengths'])]}
    assert _check_planted(hits, ns, primers, k) > 150
    a = ns['planted'][7][0][2] // 32 * 32
    assert _check_dimers(hits, dimers, wl, primers, var, 7, max(0, a - 1_000_000), a + 1_000_000) > 100


def test_config3_full_size(ctx):
    """16S pair (40 variants), k = 2, 500 Mbp in 200 records."""
    from degenprimer_b200.synthetic import make_workload
    wl = make_workload('config3_16S_500Mbp')
    tmpl = ctx.template_from_sequences([(r.ctypes.data, len(r)) for r in wl['records']])
    primers, k = wl['primers'], wl['k']
    var = _variants(primers)
    cond = _conditions()
    hits, dimers, st, dst = ctx.find(tmpl, ctx.prepare(primers, k, var, jit=True), cond=cond)
    assert st['algo'] == 3 and len(hits) > 1000
    hp, _ = ctx.scan(tmpl, primers, k, algo=1)                         # the per-window popcount kernel, all 500 Mbp
    assert np.array_equal(hits, hp)
    big = sorted(range(200), key=lambda i: -wl['lengths'][i])[:4]
    for r in big:
        _check_stretch(hits, wl['records'], wl['lengths'], primers, k, r, 0, wl['lengths'][r])
    assert _check_planted(hits, wl, primers, k) > 1000
    r = big[0]
    a = wl['planted'][r][0][2] // 32 * 32
    assert _check_dimers(hits, dimers, wl, primers, var, r, max(0, a - 200_000), a + 200_000) > 0
    tmpl.close()


@pytest.mark.parametrize('v16', [False, True])
def test_config5_sweep_full_size(ctx, v16):
    """max_mismatches 0..4 x primer length 18 / 25 / 35 on the 1 Gbp template: a 20 Mbp stretch == C oracle
    for every combination, hit counts grow with k, both kernels agree on the degenerate 35-mer."""
    from degenprimer_b200.synthetic import SWEEP_35, SWEEP_35_V16, make_workload
    wl = make_workload('config5_sweep_1Gbp')
    arr = wl['records'][0]
    tmpl = ctx.template_from_sequences([(arr.ctypes.data, len(arr))])
    base = SWEEP_35_V16 if v16 else SWEEP_35
    cond = _conditions()
    for L in (18, 25, 35):
        m = base[:L]
        last = -1
        for k in range(5):
            var = _variants([m])
            hits, dimers, st, dst = ctx.find(tmpl, ctx.prepare([m], k, var, jit=True), cond=cond)
            assert st['algo'] == 3 and len(hits) >= last
            last = len(hits)
            _check_stretch(hits, wl['records'], [len(arr)], [m], k, 0, 480_000_000, 500_000_000)
            if L == 35 and k in (2, 4):
                h2, _ = ctx.scan(tmpl, [m], k, algo=2)
                assert np.array_equal(hits, h2)
                if k == 2 and not v16:
                    assert _check_planted(hits, wl, [SWEEP_35], 2) > 60
    tmpl.close()


def test_packed_image_round_trip(ctx):
    """dp_template_save_mem / dp_template_load_mem: image in pageable and in pinned host memory gives the
    same device state and hits; a flipped bit is refused when verified."""
    import torch
    from degenprimer_b200 import _native as nat
    from tests.util import plant_sites, random_dna
    rng = np.random.default_rng(5)
    master, k = 'GAASGCRAAKATYGGGAAC', 2
    seqs = []
    for n in (50001, 33, 0, 4096, 64, 1):
        t = random_dna(rng, n, 'ACGTN', [0.2475] * 4 + [0.01])
        if n == 50001:                    # long runs: whole words, word edges, a run that ends the record
            t = t[:1000] + 'N' * 300 + t[1300:2000] + 'R' * 32 + t[2032:4064] + 'Y' * 33 + t[4097:49990] + 'N' * 11
        if n > 1000:
            t, _ = plant_sites(rng, t, [master], k, 8)
        seqs.append(t)
    t0 = ctx.template_from_sequences(seqs)
    t0.origin = 'made in a test'
    want, _ = ctx.scan(t0, [master], k)
    n = t0.image_bytes
    buf = np.zeros(n + 8, dtype=np.uint8)
    with pytest.raises(nat.NativeError):
        t0.save_image(buf.ctypes.data, n - 1)
    assert t0.save_image(buf.ctypes.data, n) == n
    pinned = torch.from_numpy(buf[:n].copy()).pin_memory()
    for ptr, verify in ((buf.ctypes.data, True), (pinned.data_ptr(), True), (pinned.data_ptr(), False)):
        t1 = ctx.template_load_image(ptr, n, verify=verify)
        assert (t1.n_records, t1.total_bases, t1.lengths, t1.n_exception_runs, t1.origin) == \
               (t0.n_records, t0.total_bases, t0.lengths, t0.n_exception_runs, 'made in a test')
        got, _ = ctx.scan(t1, [master], k)
        assert np.array_equal(got, want) and len(want) >= 8
        assert t1.unpack(0, 100, 200) == seqs[0][100:300].upper()
        # the device state is the same bit for bit -- also the exception plane, which the unverified
        # load from pinned memory does not copy but rebuilds from the record and run tables
        again = np.zeros(n, dtype=np.uint8)
        assert t1.save_image(again.ctypes.data, n) == n
        assert np.array_equal(again, buf[:n])
        t1.close()
    bad = buf[:n].copy()
    bad[n // 2] ^= 4
    with pytest.raises(nat.NativeError, match='checksum'):
        ctx.template_load_image(bad.ctypes.data, n, verify=True)
    hdr = buf[:n].copy()
    hdr[40] ^= 1                                  # total_bases: the header is under the checksum too
    with pytest.raises(nat.NativeError):
        ctx.template_load_image(hdr.ctypes.data, n, verify=True)
    with pytest.raises(nat.NativeError, match='truncated'):
        ctx.template_load_image(buf.ctypes.data, n - 8, verify=True)
    huge = buf[:n].copy()
    huge[48:56] = 0xFF                             # n_runs = 2^64 - 1: the size arithmetic must not wrap
    with pytest.raises(nat.NativeError):
        ctx.template_load_image(huge.ctypes.data, n, verify=True)


def test_cache_remembers_its_source(tmp_path):
    """the same dbname with another (older) file must not return the first file's sequences."""
    from degenprimer_b200.SeqView import SeqView
    a, b, cache = tmp_path / 'a.fa', tmp_path / 'b.fa', str(tmp_path / 'cache.dpk')
    a.write_text('>one first\nACGTACGTACGTTTGACA\n')
    b.write_text('>two second\nGGGGCCCCAAAATTTTGG\n>three\nACGT\n')
    os.utime(str(b), (1000, 1000))                                   # older than any cache written now
    v = SeqView(upper=True)
    assert v.load(str(a), dbname=cache) and v.keys() == ('one',)
    first = os.path.getmtime(cache)
    w = SeqView(upper=True)
    assert w.load(str(b), dbname=cache)
    assert w.keys() == ('three', 'two') and str(w['two'].seq) == 'GGGGCCCCAAAATTTTGG'
    x = SeqView(upper=True)
    assert x.load(str(b), dbname=cache) and x.keys() == ('three', 'two')      # now it is b's cache: reused
    os.utime(str(a), None)
    y = SeqView(upper=True)
    assert y.load(str(a), dbname=cache) and y.keys() == ('one',) and os.path.getmtime(cache) >= first


def test_find_all_equals_per_primer_find(ctx, tmp_path):
    """SearchEngine.find_all (all primers x all records, one dp_find) == find_duplexes per (record, primer);
    the lists survive pickling as SiteLists."""
    import pickle
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.SeqUtils import Seq, SeqRecord
    from degenprimer_b200.SeqView import SeqView
    from degenprimer_b200.WorkCounter import WorkCounter
    from tests.util import plant_sites, random_dna
    tdf.PCR_P.set(COND)
    rng = np.random.default_rng(21)
    masters = ['GAASGCRAAKATYGGGAAC', 'ATATTCTACRACGGCTATCC', 'ACGTTGCAAGCTTGCATGCAAGGT']
    fa = tmp_path / 'g.fa'
    with open(fa, 'w') as fh:
        for name, n in (('r1', 30011), ('r2', 64), ('r3', 12000), ('r0', 25)):
            t = random_dna(rng, n)
            if n > 1000:
                t, _ = plant_sites(rng, t, masters, 2, 6)
            fh.write('>%s test\n%s\n' % (name, t))
    view = SeqView(upper=True)
    assert view.load(str(fa))
    primers = []
    for i, m in enumerate(masters):
        p = Primer(SeqRecord(Seq(m), id='P%d' % i), 1e-6)
        p.generate_components()
        primers.append(p)
    recs = [view[k] for k in view.keys() if len(view[k]) >= 25]
    se = SearchEngine(threading.Event())
    both = se.find_all(WorkCounter(), recs, primers, 2)
    assert sorted(both) == sorted(r.id for r in recs)
    n_sites = 0
    for rec in recs:
        for i, p in enumerate(primers):
            want = se.find_duplexes(WorkCounter(), rec, p, 2)
            got = both[rec.id][i]
            if want is None:
                assert got is None or (not got[0] and not got[1])
                continue
            for strand in (0, 1):
                assert [x[0] for x in got[strand]] == [x[0] for x in want[strand]]
                back = pickle.loads(pickle.dumps(got[strand]))
                for (_, g), (_, w), (_, b) in zip(got[strand], want[strand], back):
                    assert [str(d) for d in g] == [str(d) for d in w] == [str(d) for d in b]
                    assert [[x.dG for x in d.dimers] for d in g] == [[x.dG for x in d.dimers] for d in w]
                    n_sites += 1
    assert n_sites >= 30


@pytest.mark.parametrize('world', [3, 8])
def test_sharded_genome_on_one_gpu(ctx, world):
    """distributed.ShardedGenome: the ranks' pieces scanned one after the other on this GPU, translated
    with GenomeLayout.to_global, == the whole genome (hits; and the dimers of every owned hit)."""
    from degenprimer_b200.distributed import GenomeLayout, ShardedGenome, sort_hits
    from degenprimer_b200.synthetic import make_workload
    wl = make_workload('config3_16S_500Mbp', scale=0.02)           # 10 Mbp in 200 short records
    primers, k = wl['primers'], wl['k']
    var = _variants(primers)
    cond = _conditions()
    lens = [len(p) for p in primers]
    q = ctx.prepare(primers, k, var)
    one = GenomeLayout(wl['lengths'], 1, max(lens))
    whole = ShardedGenome(ctx, one, 0, lambda r: wl['records'][r])
    h, d, _, _ = whole.find(q, cond)
    want = sort_hits(one.to_global(h, 0, lens))
    lay = GenomeLayout(wl['lengths'], world, max(lens))
    parts, n_dimers = [], 0
    for g in range(world):
        sh = ShardedGenome(ctx, lay, g, lambda r: wl['records'][r])
        hg, dg, _, _ = sh.find(q, cond)
        parts.append(lay.to_global(hg, g, lens))
        n_dimers += len(dg)
        sh.template.close()
    got = sort_hits(np.concatenate(parts))
    assert len(want) > 50 and np.array_equal(got, want) and n_dimers == len(d)


def test_matches_to_mixture_on_resident_hits_ten_conditions():
    """SURVEY §8f N2 through the facade: PCR_ProductsFinder.find_matches keeps the hits of all primers on the
    device, matches_to_mixture re-evaluates them with dp_find_again under every new set of conditions.  Ten
    trial points: annealings (positions, duplexes, every dimer's structure and dG) == the ORACLE's under the same
    conditions, and the mixture == the one the same factory builds from the oracle's annealings."""
    import json
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200.PCR_Mixture import MixtureFactory, load_mixture
    from degenprimer_b200.PCR_ProductsFinder import PCR_ProductsFinder, ResidentMatches
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqUtils import Seq, SeqRecord
    from degenprimer_b200.WorkCounter import WorkCounter
    from tests.test_reference_mixture import dump
    from tests.util import random_dna
    from oracle.sequtils import reverse_complement
    rng = np.random.default_rng(77)
    masters = [('F', 'GAASGCRAAKATYGGGAAC'), ('R', 'ATATTCTACRACGGCTATCC')]
    t = list(random_dna(rng, 40000))
    sites = []
    for i in range(10):                                   # amplicons: F site, R site (as reverse complement) 300-900 bp downstream
        at = 1500 + 3500 * i
        f = masters[0][1].replace('S', 'G').replace('R', 'A').replace('K', 'T').replace('Y', 'C')
        r = reverse_complement(masters[1][1].replace('R', 'G'))
        if i % 3 == 1:
            f = f[:4] + ('A' if f[4] != 'A' else 'C') + f[5:]          # one mismatch
        gap = 300 + 60 * i
        t[at:at + len(f)] = f
        t[at + gap:at + gap + len(r)] = r
        sites.append(at)
    text = ''.join(t)
    k = 2
    gprimers, oprimers = [], []
    for pid, m in masters:
        p = Primer(SeqRecord(Seq(m), id=pid), 1e-6)
        p.generate_components()
        gprimers.append(p)
        oprimers.append(OPrimer(OSeqRecord(m, id=pid), 1e-6))
    ppf = PCR_ProductsFinder(threading.Event(), gprimers, 50, 3000, with_exonuclease=False, include_side_annealings=True)
    mirror = MixtureFactory(threading.Event(), oprimers, 50, 3000, with_exonuclease=False, include_side_annealings=True)
    rec = SeqRecord(Seq(text), id='T', description='T synthetic')
    tdf.PCR_P.set(COND)
    matches = ppf.find_matches(WorkCounter(), rec, k)
    assert isinstance(matches, ResidentMatches) and len(matches) == 2 and matches[0][2] is gprimers[0]
    oracle_matches = [o1.find_matches(text, m, k, workers=0) for _, m in masters]
    for got, want in zip(matches, oracle_matches):
        assert np.array_equal(got[5], want[4]) and np.array_equal(got[6], want[5])
    n_dimers = 0
    for step in range(10):
        cond = dict(COND, PCR_T=52.0 + 2.0 * step, Na=(30 + 10 * step) * 1e-3, Mg=(1.0 + 0.4 * step) * 1e-3,
                    DMSO=float(step % 3))
        tdf.PCR_P.set(cond)
        ocond.PCR_P.set(cond)
        want_ann = [o1.compile_duplexes(m[0], m[1], p, m[2], m[3], m[4], m[5], mp_order=True)
                    for m, p in zip(oracle_matches, oprimers)]
        got_ann = ppf._matches_to_annealings_resident(matches)
        assert got_ann is not None and matches._state is not None                # evaluated on the resident hits
        for g, w in zip(got_ann, want_ann):
            assert (g is None) == (w is None)
            if g is None:
                continue
            for strand in (0, 1):
                assert [x[0] for x in g[strand]] == [x[0] for x in w[strand]]
                for (_, gd), (_, wd) in zip(g[strand], w[strand]):
                    assert [d.name for d in gd] == [d.name for d in wd]
                    for a, b in zip(gd, wd):
                        assert [x.fwd_matches for x in a.dimers] == [x.fwd_matches for x in b.dimers]
                        assert [x.dG for x in a.dimers] == [x.dG for x in b.dimers]          # bit for bit
                        n_dimers += len(a.dimers)
        path = ppf.matches_to_mixture(WorkCounter(), 'synthetic', matches)
        if any(w is None for w in want_ann):
            assert path is None
            continue
        fwd, rev = [], []
        for w in want_ann:
            fwd.extend(w[0])
            rev.extend(w[1])
        want_mix = mirror.create_PCR_mixture(WorkCounter(), 'synthetic', fwd, rev)
        got = json.loads(json.dumps(dump(load_mixture(path))))
        assert got == json.loads(json.dumps(dump(want_mix))) and len(got['products']) >= 5
    assert n_dimers > 200
    # hits that are gone (another search on that context) fall back to the per-primer path, same result
    ppf._resident_ctx.scan(matches._state['tmpl'], [masters[0][1]], 0)
    path = ppf.matches_to_mixture(WorkCounter(), 'synthetic', matches)
    assert matches._state is None and json.loads(json.dumps(dump(load_mixture(path)))) == got
    # a template that was closed must not be touched (use after free in dp_find_again)
    m2 = ppf.find_matches(WorkCounter(), rec, k)
    m2._state['tmpl'].close()
    with pytest.raises(Exception):
        m2._state['ctx'].find_again(tdf.conditions())
