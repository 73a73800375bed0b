"""SURVEY §8(f) N4 — the on-disk packed template (dp_template_save / dp_template_load,
SeqView.load(..., dbname=cache) / load_packed / save_packed).  CPU: the file format as the host-side
reader sees it.  -m gpu: FASTA -> pack -> save -> load gives the same device state and the same hits."""
import os
import threading

import numpy as np
import pytest

from degenprimer_b200 import _native as nat
from tests.util import plant_sites, random_dna


def _write_packed(path, records, n_runs=2, corrupt=None):
    """A packed-template file built by hand from the layout in include/degenprimer_b200.h."""
    recs = np.zeros(len(records), dtype=nat.PACK_RECORD)
    names, start = b'', 0
    for i, (rid, descr, n) in enumerate(records):
        recs[i] = (start, n, len(rid), len(descr))
        names += rid.encode() + descr.encode()
        start += (n + 31) // 32 * 32
    n_words = start // 32
    guard = 4
    runs = np.zeros(n_runs, dtype=nat.PACK_RUN)
    for i in range(n_runs):
        runs[i] = (10 * i + 3, 10 * i + 7, 14, 0)
    planes = np.arange(3 * (n_words + guard), dtype='<u4') * np.uint32(2654435761)
    h = np.zeros(1, dtype=nat.PACK_HEADER)
    h['magic'], h['version'], h['n_records'] = nat.PACK_MAGIC, 3, len(records)
    h['n_words'], h['n_words_alloc'] = n_words, n_words + guard
    h['total_bases'], h['n_runs'], h['names_bytes'] = sum(r[2] for r in records), n_runs, len(names)
    pb = planes.tobytes()
    h['checksum'] = nat.pack_checksum(h.tobytes(), recs.tobytes(), names, runs.tobytes(), pb[:len(pb) // 3],
                                      pb[len(pb) // 3:2 * len(pb) // 3], pb[2 * len(pb) // 3:])
    blob = h.tobytes() + recs.tobytes() + names + runs.tobytes() + planes.tobytes()
    if corrupt is not None:
        blob = corrupt(blob)
    with open(path, 'wb') as fh:
        fh.write(blob)
    return int(h['checksum'][0])


def test_packed_file_format_host_reader(tmp_path):
    p = str(tmp_path / 't.dpk')
    recs = [('chr1', 'chr1 first record', 1000), ('chr2', 'chr2', 33), ('empty', 'empty record', 0)]
    c = _write_packed(p, recs)
    info = nat.packed_info(p)
    assert info['n_records'] == 3 and info['total_bases'] == 1033 and info['n_runs'] == 2
    assert info['expected_bytes'] == info['bytes'] == os.path.getsize(p) and info['checksum'] == c
    assert info['records'] == [('chr1', 'chr1 first record', 0, 1000), ('chr2', 'chr2', 1024, 33),
                               ('empty', 'empty record', 1088, 0)]
    # the checksum sees every section and pads the last word of a section with zeros
    assert nat.pack_checksum(b'abc') == nat.pack_checksum(b'abc\0\0') != nat.pack_checksum(b'abd')
    _write_packed(p, recs, corrupt=lambda b: b'NOTAPACK' + b[8:])
    with pytest.raises(ValueError, match='bad magic'):
        nat.packed_info(p)
    _write_packed(p, recs, corrupt=lambda b: b[:40])
    with pytest.raises(ValueError, match='shorter'):
        nat.packed_info(p)
    _write_packed(p, recs, corrupt=lambda b: b[:-100])
    info = nat.packed_info(p)
    assert info['expected_bytes'] == info['bytes'] + 100          # truncation is visible on the host


def _fasta(path, seqs):
    with open(path, 'w') as fh:
        for name, s in seqs.items():
            fh.write('>%s synthetic %s\n' % (name, name))
            for i in range(0, len(s), 60):
                fh.write(s[i:i + 60] + '\n')


@pytest.mark.gpu
def test_packed_template_round_trip(tmp_path):
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.SeqUtils import SeqRecord, Seq
    from degenprimer_b200.SeqView import SeqView
    from degenprimer_b200.WorkCounter import WorkCounter
    master, k = 'GAASGCRAAKATYGGGAAC', 2
    rng = np.random.default_rng(8)
    seqs = {}
    for name, n, alphabet, p in (('zeta', 70001, 'ACGTN', [0.2475] * 4 + [0.01]), ('alpha', 31, 'ACGT', None),
                                 ('mid', 4096, 'ACGTRYSWKMBDHVN', [0.22] * 4 + [0.12 / 11] * 11)):
        t = random_dna(rng, n, alphabet, p)
        if n > 1000:
            t, _ = plant_sites(rng, t, [master], k, 8)
        seqs[name] = t
    seqs['zeta'] = seqs['zeta'][:30000] + 'N' * 5000 + seqs['zeta'][35000:]      # one long run
    fa, cache = str(tmp_path / 'templates.fasta'), str(tmp_path / 'templates.dpk')
    _fasta(fa, seqs)
    gp = Primer(SeqRecord(Seq(master), id='P'), 1e-6)
    gp.generate_components()
    se = SearchEngine(threading.Event())

    def hits(view):
        out = {}
        for rid in view.keys():
            rec = view[rid]
            if len(rec) < len(master):
                continue
            m = se.find_matches(WorkCounter(), rec, gp, k)
            out[rid] = (m[5].tolist(), m[6].tolist(), str(rec.seq))
        return out

    a = SeqView(upper=True)
    assert a.load(fa, dbname=cache) and os.path.isfile(cache)                      # packed from text, cache written
    want = hits(a)
    assert sum(len(v[0]) + len(v[1]) for v in want.values()) >= 8
    assert {k_: v[2] for k_, v in want.items()} == {k_: s for k_, s in seqs.items() if len(s) >= len(master)}
    info = nat.packed_info(cache)
    assert info['bytes'] == info['expected_bytes'] and info['n_records'] == 3
    assert [(r[0], r[1], r[3]) for r in info['records']] == [
        (n, '%s synthetic %s' % (n, n), len(seqs[n])) for n in sorted(seqs)]
    assert info['total_bases'] == sum(len(s) for s in seqs.values())
    raw = open(cache, 'rb').read()
    o_names = 64 + 24 * 3
    o_runs = o_names + sum(len(r[0]) + len(r[1]) for r in info['records']) + len(info['origin'].encode())
    o_planes = o_runs + 24 * info['n_runs']
    pb = (len(raw) - o_planes) // 3
    assert info['origin'].startswith(os.path.abspath(fa) + '|')
    assert nat.pack_checksum(raw[:56] + b'\0' * 8, raw[64:o_names], raw[o_names:o_runs], raw[o_runs:o_planes], raw[o_planes:o_planes + pb],
                             raw[o_planes + pb:o_planes + 2 * pb], raw[o_planes + 2 * pb:]) == info['checksum']                                                       # C writer == Python reader
    t0 = a._templates[0]
    b = SeqView(upper=True)
    assert b.load(fa, dbname=cache)                                                # cache is newer: loaded, not parsed
    t1 = b._templates[0]
    assert (t1.n_records, t1.total_bases, t1.padded_bases, t1.n_exception_runs, t1.ids, t1.descriptions,
            t1.starts, t1.lengths) == (t0.n_records, t0.total_bases, t0.padded_bases, t0.n_exception_runs, t0.ids,
                                       t0.descriptions, t0.starts, t0.lengths)
    assert hits(b) == want
    c = SeqView(upper=True)
    assert c.load_packed(cache) and hits(c) == want
    assert c.save_packed(str(tmp_path / 'again.dpk')) == [str(tmp_path / 'again.dpk')]
    assert open(tmp_path / 'again.dpk', 'rb').read() == raw                        # save(load(x)) == x, byte for byte
    # a corrupt, a truncated and a foreign file are refused
    bad = bytearray(raw)
    bad[len(bad) // 2] ^= 0x10
    for name, blob, msg in (('flip.dpk', bytes(bad), 'checksum'), ('cut.dpk', raw[:-8], 'truncated'),
                            ('magic.dpk', b'>notpack' + raw[8:], 'bad magic')):
        with open(tmp_path / name, 'wb') as fh:
            fh.write(blob)
        with pytest.raises(nat.NativeError, match=msg):
            SeqView(upper=True).load_packed(str(tmp_path / name))
    # the cache remembers size and mtime of its source: a touched FASTA file makes it stale and it is rebuilt
    os.utime(cache, (1, 1))
    d = SeqView(upper=True)
    assert d.load(fa, dbname=cache) and os.path.getmtime(cache) == 1 and hits(d) == want     # still valid: reused
    os.utime(fa, (2_000_000_000, 2_000_000_000))
    e = SeqView(upper=True)
    assert e.load(fa, dbname=cache) and os.path.getmtime(cache) > 1 and hits(e) == want


GENBANK = '''LOCUS       SCU49845     120 bp    DNA             PLN       21-JUN-1999
DEFINITION  Saccharomyces cerevisiae TCP1-beta gene, partial cds, and Axl2p
            (AXL2) and Rev7p (REV7) genes, complete cds.
ACCESSION   U49845 X00001
VERSION     U49845.1  GI:1293613
KEYWORDS    .
FEATURES             Location/Qualifiers
     source          1..120
                     /organism="Saccharomyces cerevisiae"
ORIGIN
        1 gatcctccat atacaacggt atctccacct caggtttaga tctcaacaac ggaaccattg
       61 ccgacatgag acagttaggt atcgtcgaga gttacaagct aaaacgagca gtagtcagct
//
LOCUS       NOVERSION     10 bp    DNA
DEFINITION  second record.
ACCESSION   AB000002
ORIGIN
        1 acgtnnryac
//
'''

EMBL = '''ID   X56734; SV 1; linear; mRNA; STD; PLN; 70 BP.
XX
AC   X56734; S46826;
XX
DE   Trifolium repens mRNA for non-cyanogenic beta-glucosidase
DE   second line
XX
SQ   Sequence 70 BP; 19 A; 16 C; 17 G; 18 T; 0 other;
     aaacaaacca aatatggatt ttattgtagc catatttgct ctgtttgtta ttagctcatt        60
     cacaattact                                                               70
//
ID   NOAC_NAME; linear; DNA; 8 BP.
DE   no accession
SQ   Sequence 8 BP;
     acgtacgn         8
//
'''


def test_genbank_and_embl_parsers(tmp_path):
    from degenprimer_b200.SeqView import SeqView, parse_embl, parse_genbank
    gb, em = tmp_path / 'x.gb', tmp_path / 'x.embl'
    gb.write_text(GENBANK)
    em.write_text(EMBL)
    recs = parse_genbank(str(gb))
    assert [(r[0], len(r[2])) for r in recs] == [('U49845.1', 120), ('AB000002', 10)]
    assert recs[0][1] == ('Saccharomyces cerevisiae TCP1-beta gene, partial cds, and Axl2p (AXL2) and Rev7p (REV7) '
                          'genes, complete cds.')
    assert recs[0][2].startswith('gatcctccat') and recs[0][2].endswith('gtagtcagct') and recs[1][2] == 'acgtnnryac'
    recs = parse_embl(str(em))
    assert [(r[0], len(r[2])) for r in recs] == [('X56734.1', 70), ('NOAC_NAME', 8)]
    assert recs[0][1] == 'Trifolium repens mRNA for non-cyanogenic beta-glucosidase second line'
    assert recs[0][2].endswith('cacaattact') and recs[1][2] == 'acgtacgn'
    (tmp_path / 'broken.gb').write_text(GENBANK.rsplit('//', 1)[0])
    with pytest.raises(ValueError, match='not terminated'):
        parse_genbank(str(tmp_path / 'broken.gb'))
    assert SeqView.guess_schema('a.gbk') == 'gb' and SeqView.guess_schema('a.embl') == 'embl'


@pytest.mark.gpu
def test_genbank_embl_load_and_cache(tmp_path):
    from degenprimer_b200.SeqView import SeqView
    gb, em = tmp_path / 'x.gb', tmp_path / 'x.embl'
    gb.write_text(GENBANK)
    em.write_text(EMBL)
    v = SeqView(upper=True)
    assert v.load(str(gb), dbname=str(tmp_path / 'gb.dpk'))
    assert v.keys() == ('AB000002', 'U49845.1') and str(v['AB000002'].seq) == 'ACGTNNRYAC'
    assert str(v['U49845.1'].seq).startswith('GATCCTCCAT') and len(v['U49845.1']) == 120
    assert v['U49845.1'].description.startswith('Saccharomyces cerevisiae TCP1-beta')
    info = nat.packed_info(str(tmp_path / 'gb.dpk'))
    assert [r[0] for r in info['records']] == ['AB000002', 'U49845.1']           # names travel with the cache
    w = SeqView(upper=True)
    assert w.load_packed(str(tmp_path / 'gb.dpk')) and w.keys() == v.keys()
    assert str(w['U49845.1'].seq) == str(v['U49845.1'].seq) and w['AB000002'].description == 'second record.'
    e = SeqView(upper=True)
    assert e.load(str(em)) and e.keys() == ('NOAC_NAME', 'X56734.1') and str(e['NOAC_NAME'].seq) == 'ACGTACGN'
