"""SearchEngine — the reference's call signatures over the CUDA hot path.

Drop-in for DegenPrimer/SearchEngine.py as used by
DegenPrimer/PCR_ProductsFinder.py:36,74,88,100,118-120 and re-exported at
:415-421:  `find_matches`, `compile_duplexes`, `compile_duplexes_mp`, `find`,
`batch_find`, `mp_better`, `set_max_chunk`, module-level `find` / `batch_find`.
The FFT correlation and the per-site Python Duplex construction are replaced by
dp_scan / dp_duplex_dg (include/degenprimer_b200.h); return values keep the
reference's shapes and conventions:

  find_matches -> (fwd_seq, rev_seq, primer, t_len, p_len, fwd_matches, rev_matches)
                  rev_matches index the reverse-complement strand
  compile_*    -> ([(pos, duplexes), ...], [(pos, duplexes), ...]) or None,
                  pos = 1-based forward coordinate of the primer's 3' end
  find         -> path of a temp store holding that pair under 'result'
"""
import os
from collections import OrderedDict

import numpy as np

from . import TD_Functions as tdf
from . import _native as nat
from .SecStructures import DuplexTable, LazyWindows, SiteDuplexes, SiteList
from .SeqView import ResidentRecord, StrandView
from .WorkCounter import WorkCounter
from .tmpStorage import to_shelf

cpu_count = os.cpu_count() or 1


class DummyEvent(object):
    """BioUtils/Tools/AbortableBase.py:11-14."""
    def is_set(self):
        return False

    def set(self):
        pass


class _PackCache(object):
    """Templates packed from plain strings, keyed by object identity, so that a
    caller who hands the same 7-tuple back many times (PCR_Optimizer re-runs the
    dG stage per optimiser step, DegenPrimer/PCR_Optimizer.py:162-169) uploads once."""

    def __init__(self, size=4):
        self._d = OrderedDict()
        self._size = size

    def get(self, obj):
        hit = self._d.get(id(obj))
        if hit is not None and hit[0] is obj:
            self._d.move_to_end(id(obj))
            return hit[1]
        return None

    def put(self, obj, template):
        self._d[id(obj)] = (obj, template)
        while len(self._d) > self._size:
            self._d.popitem(last=False)


_pack_cache = _PackCache()


def resident(template):
    """(native Template, record index) for any template-like argument."""
    if isinstance(template, ResidentRecord):
        return template._template, template._record
    seq = getattr(template, 'seq', template)
    if isinstance(seq, StrandView):
        if seq.reverse:
            raise ValueError('expected the forward strand of a resident record')
        return seq.template, seq.record
    cached = _pack_cache.get(seq)
    if cached is None:
        text = str(seq)
        cached = tdf._ctx().template_from_sequences([text])
        cached.host_text = text
        _pack_cache.put(seq, cached)
    return cached, 0


class SearchEngine(object):
    """Mismatch-tolerant search of degenerate primers in template sequences and
    evaluation of the duplexes they form, on the GPU."""

    _max_chunk_size = 2 ** 12     # the reference's FFT chunk; meaningless here, kept for set_max_chunk

    def __init__(self, abort_event=None, jit=None):
        """jit: compile the scan kernel for every primer set this engine meets (NVRTC, ~0.3 s once
        per set and process; pays for Gbp templates and repeated searches).  None: the
        DEGENPRIMER_B200_JIT environment variable decides (default off)."""
        self._abort_event = abort_event if abort_event is not None else DummyEvent()
        self._jit = jit
        self.last_stats = {}

    # ---- plumbing the reference exposes ----------------------------------------------
    @classmethod
    def set_max_chunk(cls, chunk_size):
        cls._max_chunk_size = chunk_size

    def aborted(self):
        return self._abort_event.is_set()

    @staticmethod
    def _check_length_inequality(t_len, p_len):
        if t_len < p_len or p_len == 0:
            raise ValueError('SearchEngine._find: template sequence should be '
                             'longer or equal to primer sequence and both '
                             'should be grater than zero.')

    @staticmethod
    def mp_better(t_len):
        """DegenPrimer/SearchEngine.py:241-245; iPCR routes templates with it."""
        return cpu_count > 1 and t_len > 25000

    # ---- scan -------------------------------------------------------------------------
    def _scan(self, tmpl, rec, masters, mismatches):
        ctx = tmpl.ctx
        begin = tmpl.starts[rec]
        end = begin + ((tmpl.lengths[rec] + 31) // 32) * 32
        if tmpl.n_records == 1:
            begin, end = 0, 2 ** 64 - 1
        hits, stats = ctx.scan(tmpl, masters, mismatches, begin, end)
        self.last_stats['scan'] = stats
        return hits

    def find_matches(self, counter, template, primer, mismatches):
        """All windows of both strands whose thirds-weighted score reaches
        max(1, L-k) - 0.5 (DegenPrimer/SearchEngine.py:362-372 -> :280-359)."""
        p_len, t_len = len(primer), len(template)
        self._check_length_inequality(t_len, p_len)
        if self.aborted():
            return None
        tmpl, rec = resident(template)
        counter.set_work(t_len)
        try:
            hits = self._scan(tmpl, rec, [str(primer.master_sequence.seq)], int(mismatches))
        except nat.Aborted:
            return None
        counter.done()
        if self.aborted():
            return None
        hits = hits[hits['record'] == rec]
        fwd_matches = hits['pos'][hits['strand'] == 0].astype(np.int64)
        rev_matches = hits['pos'][hits['strand'] == 1].astype(np.int64)
        text = getattr(tmpl, 'host_text', None)
        fwd_seq = StrandView(tmpl, rec, False, text)
        rev_seq = StrandView(tmpl, rec, True)
        return fwd_seq, rev_seq, primer, t_len, p_len, fwd_matches, rev_matches

    # ---- duplex thermodynamics ------------------------------------------------------------
    @staticmethod
    def _site_results(tmpl, hits, dimers, primer, t_len, p_len, variants, names, T, n_f):
        """(fwd_results, rev_results): lists of (position, SiteDuplexes) in hit order,
        sites without a stable duplex dropped (DegenPrimer/SearchEngine.py:106-126).
        hits: forward hits first (n_f of them), then reverse; dimers grouped by hit."""
        table = _LazyWindowsTable(dimers, tmpl, hits, p_len, variants, names, T)
        bounds = np.searchsorted(dimers['hit'], np.arange(len(hits) + 1))
        pos = hits['pos'].astype(np.int64)
        site_pos = np.where(hits['strand'] == 0, pos + p_len, t_len + 1 - (pos + p_len))
        fwd_results, rev_results = [], []
        for h in np.flatnonzero(bounds[1:] > bounds[:-1]):
            item = (int(site_pos[h]), SiteDuplexes(table, int(h), int(bounds[h]), int(bounds[h + 1])))
            (fwd_results if h < n_f else rev_results).append(item)
        return fwd_results, rev_results

    def _compile(self, fwd_seq, rev_seq, primer, t_len, p_len, fwd_matches, rev_matches):
        """Both strands in one dp_duplex_dg call."""
        if isinstance(fwd_seq, StrandView):
            tmpl, rec = fwd_seq.template, fwd_seq.record
        else:
            tmpl, rec = resident(fwd_seq)
        n_f, n_r = len(fwd_matches), len(rev_matches)
        hits = np.zeros(n_f + n_r, dtype=nat.HIT_DTYPE)
        hits['pos'][:n_f] = fwd_matches
        hits['pos'][n_f:] = rev_matches
        hits['strand'][n_f:] = 1
        hits['record'] = rec
        records = primer.seq_records
        variants = [str(r.seq) for r in records]
        names = [r.id for r in records]
        with tdf.AcquireParameters():       # the snapshot is the by-value conditions block
            cond = tdf.conditions()
            T = tdf.PCR_P.PCR_T
        dimers, stats = tmpl.ctx.duplex_dg(tmpl, hits, [p_len], [len(variants)], [0],
                                           ''.join(variants), cond)
        self.last_stats['dg'] = stats
        return self._site_results(tmpl, hits, dimers, primer, t_len, p_len, variants, names, T, n_f)

    def _find_fused(self, template, primer, mismatches):
        """find_matches + compile_duplexes_mp as ONE stream submission (dp_find):
        the hits never leave the device between the two stages."""
        p_len, t_len = len(primer), len(template)
        tmpl, rec = resident(template)
        begin = tmpl.starts[rec]
        end = begin + ((tmpl.lengths[rec] + 31) // 32) * 32
        if tmpl.n_records == 1:
            begin, end = 0, 2 ** 64 - 1
        query, variants, names = _primer_query(primer, int(mismatches))
        with tdf.AcquireParameters():       # the snapshot is the by-value conditions block
            cond = tdf.conditions()
            T = tdf.PCR_P.PCR_T
        hits, dimers, st, dst = tmpl.ctx.find(tmpl, query, cond=cond, seg_begin=begin, seg_end=end)
        self.last_stats['scan'] = st
        self.last_stats['dg'] = dst
        if len(hits) and not (hits['record'] == rec).all():
            raise RuntimeError('internal: hits outside the scanned record')
        if not len(hits):
            return None
        n_f = int(np.count_nonzero(hits['strand'] == 0))
        fwd_results, rev_results = self._site_results(tmpl, hits, dimers, primer, t_len, p_len,
                                                      variants, names, T, n_f)
        if not fwd_results and not rev_results:
            return None
        fwd_results.sort(key=lambda x: x[0])
        rev_results.sort(key=lambda x: x[0])
        return fwd_results, rev_results

    def compile_duplexes_mp(self, counter, fwd_seq, rev_seq, primer, t_len, p_len,
                            fwd_matches, rev_matches, num_jobs=None):
        """DegenPrimer/SearchEngine.py:129-157: both lists sorted by position."""
        if not len(fwd_matches) + len(rev_matches):
            return None
        counter.set_subwork(2, [max(1, len(fwd_matches)), max(1, len(rev_matches))])
        if self.aborted():
            return None
        try:
            fwd_results, rev_results = self._compile(fwd_seq, rev_seq, primer, t_len, p_len,
                                                     fwd_matches, rev_matches)
        except nat.Aborted:
            return None
        counter[0].done()
        counter[1].done()
        if self.aborted() or not fwd_results and not rev_results:
            return None
        fwd_results.sort(key=lambda x: x[0])
        rev_results.sort(key=lambda x: x[0])
        return fwd_results, rev_results

    def compile_duplexes(self, counter, fwd_seq, rev_seq, primer, t_len, p_len,
                         fwd_matches, rev_matches):
        """DegenPrimer/SearchEngine.py:160-187: match order, empty lists become None."""
        if not len(fwd_matches) + len(rev_matches):
            return None
        counter.set_work(len(fwd_matches) + len(rev_matches))
        if self.aborted():
            return None
        try:
            fwd_results, rev_results = self._compile(fwd_seq, rev_seq, primer, t_len, p_len,
                                                     fwd_matches, rev_matches)
        except nat.Aborted:
            return None
        counter.done()
        if self.aborted():
            return None
        if not fwd_results:
            fwd_results = None
        if not rev_results:
            rev_results = None
        if not fwd_results and not rev_results:
            return None
        return fwd_results, rev_results

    # ---- compositions -------------------------------------------------------------------------
    def find_duplexes(self, counter, template, primer, mismatches):
        """find_matches + compile_duplexes_mp (what `find` puts on the shelf), fused on
        the device; returns (fwd_results, rev_results) or None."""
        p_len, t_len = len(primer), len(template)
        self._check_length_inequality(t_len, p_len)
        counter.set_subwork(2, (t_len, t_len * primer.num_components))
        if self.aborted():
            return None
        try:
            duplexes = self._find_fused(template, primer, mismatches)
        except nat.Aborted:
            return None
        counter[0].done()
        counter[1].done()
        if self.aborted():
            return None
        return duplexes

    def find(self, counter, template, primer, mismatches):
        """DegenPrimer/SearchEngine.py:375-386."""
        duplexes = self.find_duplexes(counter, template, primer, mismatches)
        return to_shelf(duplexes) if duplexes else None

    # ---- all primers x all records of a resident template in ONE device pass --------------------
    def find_all(self, counter, templates, primers, mismatches):
        """`find` for every (template, primer) pair at once: the templates must be records of one
        resident template (a loaded SeqView, ResidentRecords); the scan of all primers over all of
        them and the duplex stage of every site are ONE dp_find submission instead of the
        per-primer, per-template loop of DegenPrimer/PCR_ProductsFinder.py:94-111,114-141.
        Returns {template.id: [(fwd_results, rev_results) or None, one per primer]} with the lists
        of `find_duplexes` (sorted by position), or None when aborted."""
        templates = list(templates)
        primers = list(primers)
        if not templates or not primers:
            return {}
        native = [resident(t) for t in templates]
        tmpl = native[0][0]
        if any(n[0] is not tmpl for n in native):
            raise ValueError('find_all: the templates must be records of one resident template')
        recs = [n[1] for n in native]
        for t in templates:
            for p in primers:
                self._check_length_inequality(len(t), len(p))
        counter.set_work(len(templates) * len(primers))
        if self.aborted():
            return None
        query, var_lists, name_lists = _primers_query(tmpl.ctx, primers, int(mismatches), self._jit)
        begin = min(tmpl.starts[r] for r in recs)
        end = max(tmpl.starts[r] + ((tmpl.lengths[r] + 31) // 32) * 32 for r in recs)
        if len(set(recs)) == tmpl.n_records:
            begin, end = 0, 2 ** 64 - 1
        with tdf.AcquireParameters():       # the snapshot is the by-value conditions block
            cond = tdf.conditions()
            T = tdf.PCR_P.PCR_T
        try:
            hits, dimers, st, dst = tmpl.ctx.find(tmpl, query, cond=cond, seg_begin=begin, seg_end=end)
        except nat.Aborted:
            return None
        self.last_stats['scan'] = st
        self.last_stats['dg'] = dst
        if self.aborted():
            return None
        out = self._split_results(tmpl, templates, recs, primers, hits, dimers, var_lists, name_lists, T)
        counter.done()
        return out

    @staticmethod
    def _split_results(tmpl, templates, recs, primers, hits, dimers, var_lists, name_lists, T):
        """hits sorted by (primer, strand, record, pos), dimers grouped by hit -> per (template,
        primer) a pair of SiteLists; no per-site Python work."""
        n_p = len(primers)
        p_lens = np.array([len(p) for p in primers], dtype=np.int64)
        windows = LazyWindows(tmpl, hits, p_lens)
        tables = [DuplexTable(dimers, windows, var_lists[i], name_lists[i], T) for i in range(n_p)]
        out = {t.id: [None] * n_p for t in templates}
        if not len(hits):
            return out
        bounds = np.searchsorted(dimers['hit'], np.arange(len(hits) + 1))
        keep = np.flatnonzero(bounds[1:] > bounds[:-1])            # sites with a stable duplex
        if not len(keep):
            return out
        kh = hits[keep]
        L = p_lens[kh['primer']]
        pos = kh['pos'].astype(np.int64)
        t_len = np.asarray(tmpl.lengths, dtype=np.int64)[kh['record']]
        site = np.where(kh['strand'] == 0, pos + L, t_len + 1 - (pos + L))
        lo, hi = bounds[keep], bounds[keep + 1]
        n_r = tmpl.n_records
        key = (kh['primer'].astype(np.int64) * 2 + kh['strand']) * n_r + kh['record']
        cut = np.flatnonzero(np.diff(key)) + 1
        starts = np.concatenate(([0], cut))
        ends = np.concatenate((cut, [len(key)]))
        by_rec = {}
        for t, r in zip(templates, recs):
            by_rec.setdefault(r, []).append(t.id)
        for a, b in zip(starts, ends):
            k0 = int(key[a])
            r = k0 % n_r
            ids = by_rec.get(r)
            if not ids:
                continue
            p_i, strand = (k0 // n_r) // 2, (k0 // n_r) % 2
            sl = slice(a, b) if strand == 0 else slice(b - 1, a - 1 if a else None, -1)   # ascending position
            lst = SiteList(tables[p_i], keep[sl], site[sl], lo[sl], hi[sl])
            for tid in ids:
                pair = out[tid][p_i]
                if pair is None:
                    pair = out[tid][p_i] = [SiteList(tables[p_i], keep[:0], site[:0], lo[:0], hi[:0]),
                                            SiteList(tables[p_i], keep[:0], site[:0], lo[:0], hi[:0])]
                pair[strand] = lst
        for tid, pairs in out.items():
            out[tid] = [tuple(p) if p is not None else None for p in pairs]
        return out

    def _batch_find(self, template, primer, p_len, mismatches):
        """DegenPrimer/SearchEngine.py:388-394 (serial orderings)."""
        self._check_length_inequality(len(template), p_len)
        matches = self.find_matches(WorkCounter(), template, primer, mismatches)
        if matches is None:
            return None
        duplexes = self.compile_duplexes(WorkCounter(), *matches)
        return (template.id, to_shelf(duplexes)) if duplexes else None

    def batch_find(self, counter, templates, primer, mismatches, **kwargs):
        """DegenPrimer/SearchEngine.py:396-411: {template.id: result path}."""
        templates = list(templates)
        if not templates:
            return {}
        counter.set_work(len(templates))
        results = {}
        for template in templates:
            if self.aborted():
                return None
            result = self._batch_find(template, primer, len(primer), mismatches)
            if result:
                results[result[0]] = result[1]
            counter.count()
        return results


def _primer_query(primer, mismatches):
    """(native Query, variant strings, variant names) of a primer, cached on the primer
    object: the reference calls `find` once per template with the same Primer."""
    cache = getattr(primer, '_dp_query_cache', None)
    master = str(primer.master_sequence.seq)
    key = (master, mismatches, primer.num_components)
    if cache is not None and cache[0] == key:
        return cache[1]
    records = primer.seq_records
    variants = [str(r.seq) for r in records]
    names = [r.id for r in records]
    value = (nat.Query([master], mismatches, [variants]), variants, names)
    try:
        primer._dp_query_cache = (key, value)
    except AttributeError:          # a foreign Primer class with __slots__
        pass
    return value


def _primers_query(ctx, primers, mismatches, jit=None):
    """(native Query over all primers, variant strings per primer, variant names per primer),
    cached on the first primer object by the masters of the set."""
    masters = [str(p.master_sequence.seq) for p in primers]
    key = (tuple(masters), mismatches, tuple(p.num_components for p in primers), bool(jit))
    cache = getattr(primers[0], '_dp_set_cache', None)
    if cache is not None and cache[0] == key:
        return cache[1]
    var_lists = [[str(r.seq) for r in p.seq_records] for p in primers]
    name_lists = [[r.id for r in p.seq_records] for p in primers]
    value = (ctx.prepare(masters, mismatches, var_lists, jit=jit), var_lists, name_lists)
    try:
        primers[0]._dp_set_cache = (key, value)
    except AttributeError:
        pass
    return value


class _LazyWindowsTable(DuplexTable):
    """DuplexTable whose window strings are fetched from HBM on first use."""

    def __init__(self, dimers, tmpl, hits, p_len, variants, names, T):
        DuplexTable.__init__(self, dimers, None, variants, names, T)
        self._tmpl, self._hits, self._p_len = tmpl, hits, p_len

    @property
    def windows(self):
        if self._windows is None:
            self._windows = self._tmpl.ctx.hit_windows(self._tmpl, self._hits, [self._p_len])
            self._tmpl = None
        return self._windows

    @windows.setter
    def windows(self, value):
        self._windows = value


# shortcut functions (DegenPrimer/SearchEngine.py:415-421)
mp_better = SearchEngine.mp_better


def find(counter, template, primer, mismatches, abort_event):
    return SearchEngine(abort_event).find(counter, template, primer, mismatches)


def batch_find(counter, templates, primer, mismatches, abort_event, **kwargs):
    return SearchEngine(abort_event).batch_find(counter, templates, primer, mismatches, **kwargs)
