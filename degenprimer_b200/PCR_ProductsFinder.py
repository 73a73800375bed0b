"""PCR_ProductsFinder — host mirror of DegenPrimer/PCR_ProductsFinder.py:31-141: the caller of the hot
path.  For every primer of the system it runs the scan + duplex stage (SearchEngine, i.e. the CUDA path)
on a template, pools the annealing sites of all primers and pairs them into products (MixtureFactory).
Same constructor and method names, argument order and return values (paths of saved mixtures, `None`
when aborted or when nothing was found); the reference's Manager wrappers (:144-168) are not needed — there
is no process fan-out to coordinate, the GPU grid is the fan-out.
"""
import numpy as np

from . import TD_Functions as tdf
from . import _native as nat
from .PCR_Mixture import MixtureFactory
from .SearchEngine import SearchEngine, StrandView, _primers_query, resident
from .SecStructures import SiteList, SitePool
from .SeqUtils import pretty_rec_name
from .WorkCounter import WorkCounter
from .tmpStorage import cleanup_files, from_shelf


class ResidentMatches(list):
    """What `find_matches` returns: the reference's list of 7-tuples, one per primer
    (DegenPrimer/PCR_ProductsFinder.py:83-92), plus a handle on the same hits still sitting in HBM on a
    context of their own.  `matches_to_mixture` re-evaluates them there under the conditions of the
    moment (dp_find_again) — the optimiser calls it once per trial point
    (DegenPrimer/PCR_Optimizer.py:162-214) — instead of uploading the positions again per primer and step."""

    def __init__(self, tuples, state):
        list.__init__(self, tuples)
        self._state = state          # dict(ctx, tmpl, rec, template, hits, var_lists, name_lists) or None


class PCR_ProductsFinder(MixtureFactory):
    def __init__(self, abort_event, *args, **kwargs):
        MixtureFactory.__init__(self, abort_event, *args, **kwargs)
        self._searcher = SearchEngine(self._abort_event)
        self._resident_ctx = None
        self._p_weights = [p.num_components for p in self._primers]
        self._pw_sum = sum(self._p_weights)
        self._num_p = len(self._primers)

    @property
    def searcher(self):
        return self._searcher

    @staticmethod
    def _pool(results):
        """One forward and one reverse list out of the (fwd_results, rev_results) pairs of several primers."""
        results = list(results)
        lists = [pair[strand] for pair in results if pair for strand in (0, 1) if pair[strand]]
        if lists and all(isinstance(x, SiteList) for x in lists):
            pooled = (SitePool(), SitePool())          # device results stay arrays: MixtureFactory joins them on the GPU
        else:
            pooled = ([], [])
        for pair in results:
            for strand in (0, 1):
                if pair and pair[strand]:
                    pooled[strand].extend(pair[strand])
        return pooled

    @classmethod
    def _combine_annealings(cls, *annealings_list):
        """Results held in memory (PCR_ProductsFinder.py:57-65)."""
        return cls._pool(annealings_list)

    @classmethod
    def _extract_annealings(cls, *paths):
        """Results handed over as temp files (:45-55); an empty store contributes nothing."""
        return cls._pool(from_shelf(path) for path in paths)

    def find_matches(self, counter, template, mismatches):
        """Scan only: one 7-tuple per primer (:83-92).  All primers are ONE device pass; the hits stay
        resident for `matches_to_mixture` (ResidentMatches)."""
        if isinstance(self._searcher, SearchEngine):
            return self._find_matches_resident(counter, template, mismatches)
        return self._find_matches_per_primer(counter, template, mismatches)

    def _find_matches_resident(self, counter, template, mismatches):
        for primer in self._primers:
            self._searcher._check_length_inequality(len(template), len(primer))
        counter.set_work(self._pw_sum)
        if self.aborted():
            return None
        tmpl, rec = resident(template)
        if self._resident_ctx is None or self._resident_ctx.device != tmpl.ctx.device:
            self._resident_ctx = tdf.new_context(tmpl.ctx.device)     # nobody else's calls drop these hits
        ctx = self._resident_ctx
        query, var_lists, name_lists = _primers_query(ctx, self._primers, int(mismatches), self._searcher._jit)
        begin = tmpl.starts[rec]
        end = begin + ((tmpl.lengths[rec] + 31) // 32) * 32
        if tmpl.n_records == 1:
            begin, end = 0, 2 ** 64 - 1
        with tdf.AcquireParameters():
            cond = tdf.conditions()
        try:
            hits, _dimers, st, dst = ctx.find(tmpl, query, cond=cond, seg_begin=begin, seg_end=end)
        except nat.Aborted:
            return None
        self._searcher.last_stats['scan'], self._searcher.last_stats['dg'] = st, dst
        if self.aborted():
            return None
        counter.done()
        t_len = len(template)
        text = getattr(tmpl, 'host_text', None)
        fwd_seq, rev_seq = StrandView(tmpl, rec, False, text), StrandView(tmpl, rec, True)
        tuples = []
        for i, primer in enumerate(self._primers):
            h = hits[(hits['primer'] == i) & (hits['record'] == rec)]
            tuples.append((fwd_seq, rev_seq, primer, t_len, len(primer),
                           h['pos'][h['strand'] == 0].astype(np.int64), h['pos'][h['strand'] == 1].astype(np.int64)))
        return ResidentMatches(tuples, dict(ctx=ctx, tmpl=tmpl, rec=rec, template=template, hits=hits,
                                            var_lists=var_lists, name_lists=name_lists))

    def _matches_to_annealings_resident(self, matches_list):
        """(fwd, rev) annealing lists of every primer from the resident hits, or None when they are gone
        (another search used the context, the template was closed): the caller then goes per primer."""
        st = getattr(matches_list, '_state', None)
        if st is None or len(matches_list) != self._num_p:
            return None
        with tdf.AcquireParameters():
            cond = tdf.conditions()
            T = tdf.PCR_P.PCR_T
        try:
            dimers, dst = st['ctx'].find_again(cond)
        except nat.NativeError as e:
            if e.code != nat.DP_ERR_STATE:
                raise
            matches_list._state = None
            return None
        self._searcher.last_stats['dg'] = dst
        found = SearchEngine._split_results(st['tmpl'], [st['template']], [st['rec']], self._primers, st['hits'],
                                            dimers, st['var_lists'], st['name_lists'], T)
        return found[st['template'].id]          # one entry per primer; None = no stable duplex for that primer

    def _find_matches_per_primer(self, counter, template, mismatches):
        all_matches = []
        counter.set_subwork(self._num_p, self._p_weights)
        for i, primer in enumerate(self._primers):
            if self.aborted():
                return None
            matches = self._searcher.find_matches(counter[i], template, primer, mismatches)
            if matches is None:
                return None
            all_matches.append(matches)
        return all_matches

    def matches_to_mixture(self, counter, tname, matches_list):
        """Duplex stage + pairing for matches found earlier — the optimiser's inner loop (:67-81).  Matches
        that came from `find_matches` are re-evaluated where they are, on the device (dp_find_again)."""
        all_annealings = []
        weights = [len(m) for m in matches_list]
        counter.set_subwork(len(matches_list) + 1, weights + [0.01 * sum(weights)])
        if self.aborted():
            return None
        resident_annealings = self._matches_to_annealings_resident(matches_list)
        if resident_annealings is not None:
            if any(pair is None for pair in resident_annealings):
                return None              # the reference gives up when one primer has no annealing (:74-75)
            mixture = self.create_PCR_mixture(counter[-1], tname, *self._combine_annealings(*resident_annealings))
            counter.done()
            return mixture.save()
        for i, matches in enumerate(matches_list):
            if self.aborted():
                return None
            annealings = self._searcher.compile_duplexes_mp(counter[i], *matches)
            if annealings is None:
                return None
            all_annealings.append(annealings)
        mixture = self.create_PCR_mixture(counter[-1], tname, *self._combine_annealings(*all_annealings))
        counter.done()
        return mixture.save()          # like the reference, this raises when no product was found

    def find(self, counter, template, mismatches):
        """{template name: path of the saved mixture}, or None (:94-112).  The reference runs
        `SearchEngine.find` once per primer (one template pass and one temp file each); here all
        primers of the system are ONE device pass (SearchEngine.find_all) and the annealings stay in
        memory."""
        counter.set_subwork(2, [self._pw_sum, 0.01 * self._pw_sum])
        if self.aborted():
            return None
        found = self._searcher.find_all(counter[0], [template], self._primers, mismatches)
        if found is None:
            return None
        all_annealings = [pair for pair in found[template.id] if pair is not None]
        if not all_annealings:
            return None
        tname = pretty_rec_name(template)
        mixture = self.create_PCR_mixture(counter[-1], tname, *self._combine_annealings(*all_annealings))
        counter.done()
        if not mixture:
            return None
        return {tname: mixture.save()}

    def batch_find(self, counter, templates, mismatches, **kwargs):
        """Many short templates: {template name: path}, or None when nothing was found (:114-141)."""
        results = dict()
        counter.set_subwork(1)
        counter[0].set_work(max(1, len(templates)))
        for template in templates:
            if self.aborted():
                return None
            search_results = []
            for primer in self._primers:
                matches = self._searcher.find_matches(WorkCounter(), template, primer, mismatches)
                if not matches:
                    continue
                duplexes = self._searcher.compile_duplexes(WorkCounter(), *matches)
                if duplexes:
                    search_results.append(duplexes)
            counter[0].count()
            if not search_results:
                continue
            fwd, rev = self._combine_annealings(*search_results)
            tname = pretty_rec_name(template)
            mixture = self.create_PCR_mixture(WorkCounter(), tname, fwd, rev)
            if mixture:
                results[tname] = mixture.save()
        return results or None
