"""PCR_Mixture / MixtureFactory — host mirror of DegenPrimer/PCR_Mixture.py:29-249.

The step right after the hot path: annealing sites of both strands (the `(position, duplexes)` lists
SearchEngine.compile_duplexes[_mp] returns) are split into sites that can prime a product and sites
that cannot, and forward / reverse sites are paired into products of [min_amplicon, max_amplicon].
This is host code on purpose: the pairing loop of the reference carries state from one forward site
to the next (see `create_PCR_mixture`), so its exact result is a sequential walk over two sorted lists,
and nearly all of its time is object bookkeeping that PCR_Simulation reads (`reaction_id, annealings,
templates, products`): 674 sites -> 3393 products take 139 ms here, 173 ms in the reference (same box).
"""
import numpy as np

from . import tmpStorage
from .Product import Product, Region
from .SecStructures import SitePool


class PCR_Mixture(object):
    """All annealings, template regions and products of one target sequence (PCR_Mixture.py:29-107)."""
    __slots__ = ['reaction_id', 'annealings', 'templates', 'products']

    def __init__(self, reaction_id):
        self.reaction_id = reaction_id
        self.annealings = []            # (duplexes, template Region) of every site that takes part
        self.templates = [[], []]       # [reverse strand, forward strand]: merged template regions
        self.products = dict()          # hash(Product) -> Product

    def _compact_templates(self, strand):
        """Merge overlapping regions of a start-sorted list, in place on the Region objects (:62-69) —
        the same objects sit in `annealings`, which therefore see the merged extent too."""
        regions = self.templates[strand]
        merged = [regions[0]]
        for r in regions:
            if r.overlaps(merged[-1]):
                merged[-1] += r
            else:
                merged.append(r)
        self.templates[strand] = merged

    def add_annealing(self, duplexes, template):
        self.annealings.append((duplexes, template))
        strand = template.forward
        self.templates[strand].append(template)
        self.templates[strand].sort(key=lambda x: x.start)
        self._compact_templates(strand)

    def add_annealings(self, annealings):
        self.annealings.extend(annealings)
        before = (len(self.templates[0]), len(self.templates[1]))
        for _dups, template in annealings:
            self.templates[template.forward].append(template)
        for strand in (1, 0):
            if len(self.templates[strand]) > before[strand]:
                self.templates[strand].sort(key=lambda x: x.start)
                self._compact_templates(strand)

    def add_product(self, start, end, fwd_duplexes, rev_duplexes):
        """start, end: 1-based positions of the first and last base of the product (:72-92)."""
        new = Product(self.reaction_id, start, end, fwd_duplexes, rev_duplexes)
        key = hash(new)
        if key in self.products:
            self.products[key] += new
        else:
            self.products[key] = new

    def save(self):
        """Temp-file hand-off like the reference's tmpDict (:95-106); read back with `load_mixture`."""
        return tmpStorage.to_shelf(self)

    def __getstate__(self):
        return tuple(getattr(self, s) for s in self.__slots__)

    def __setstate__(self, state):
        for s, v in zip(self.__slots__, state):
            setattr(self, s, v)


class LazyMixture(object):
    """The PCR_Mixture of a device join, kept as arrays until somebody looks into it.

    MixtureFactory builds it from SitePools (the structure-of-arrays annealing lists of the CUDA path): the
    good / bad 3' split is a few numpy reductions over the dimer rows, the pairing loop is dp_pair_products
    on the GPU.  What exists then is the list of (forward site, reverse site) products; the objects
    PCR_Simulation reads -- `annealings`, `templates`, `products` with their Duplex, Region and Product
    instances -- are made on first access by replaying the reference's calls (add_product / add_annealing,
    DegenPrimer/PCR_Mixture.py:216-249) in the reference's order over that list, so they are what the
    reference builds, aliasing included.  Pickles as arrays (`save()`), materialises wherever it is loaded."""

    def __init__(self, reaction_id, state):
        self.reaction_id = reaction_id
        self._state = state
        self._real = None

    # ---- what exists without materialising --------------------------------------------------------
    @property
    def n_products(self):
        """Number of (forward site, reverse site) pairs; products that coincide are merged on access."""
        return len(self._state['pairs'])

    def product_bounds(self):
        """(start, end) of every pair as two int64 arrays, in the order the reference finds them."""
        st = self._state
        return (st['good_pos'][1][st['pairs']['fwd']] + 1, st['good_pos'][0][st['pairs']['rev']] - 1)

    # ---- the reference's attributes -----------------------------------------------------------------
    annealings = property(lambda self: self._materialise().annealings)
    templates = property(lambda self: self._materialise().templates)
    products = property(lambda self: self._materialise().products)

    def save(self):
        return tmpStorage.to_shelf(self)

    def __reduce__(self):
        return (LazyMixture, (self.reaction_id, self._state))

    def _entry(self, cache, kind, strand, k):
        """(position, duplexes that count, template Region) of good site k / bad site k of a strand."""
        st = self._state
        key = (kind, strand, k)
        hit = cache.get(key)
        if hit is not None:
            return hit
        part_i, site_i = st[kind + '_part'][strand][k], st[kind + '_site'][strand][k]
        pool = st['pools'][strand]
        part = pool.parts[part_i]
        pos, duplexes = part[int(site_i)]
        rkey = ('region', strand, int(part_i), int(site_i))
        region = cache.get(rkey)
        if region is None:
            length = len(part._table.variants[0])
            if strand:
                region = Region(self.reaction_id, pos - length + 1, pos, strand)
            else:
                region = Region(self.reaction_id, pos, pos + length - 1, strand)
            cache[rkey] = region
        flags = st['dup_good'][strand][part_i]
        bounds = st['dup_bounds'][strand][part_i]
        known = st['dup_known'][strand][part_i]
        a, b = int(bounds[site_i]), int(bounds[site_i + 1])
        want = (flags[a:b] & known[a:b]) if kind == 'good' else (~flags[a:b] & known[a:b])
        dups = [d for d, ok in zip(duplexes, want) if ok]
        out = cache[key] = (int(pos), dups, region)
        return out

    def _materialise(self):
        if self._real is not None:
            return self._real
        st = self._state
        mixture = PCR_Mixture(self.reaction_id)
        cache = {}
        added = (set(), set())
        rs = st['rev_start']
        for f, j in zip(st['pairs']['fwd'].tolist(), st['pairs']['rev'].tolist()):
            fwd_pos, fwd_dups, fwd_templ = self._entry(cache, 'good', 1, f)
            rev_pos, rev_dups, rev_templ = self._entry(cache, 'good', 0, j)
            mixture.add_product(fwd_pos + 1, rev_pos - 1, fwd_dups, rev_dups)
            if f not in added[1]:
                mixture.add_annealing(fwd_dups, fwd_templ)
                added[1].add(f)
            rel = j - int(rs[f])                      # the reference remembers reverse sites by their index in the slice
            if rel not in added[0]:
                mixture.add_annealing(rev_dups, rev_templ)
                added[0].add(rel)
        if st['include_side_annealings']:
            for strand in (1, 0):
                side = [self._entry(cache, 'good', strand, k) for k in range(len(st['good_pos'][strand]))
                        if k not in added[strand]]
                MixtureFactory._add_side_annealings(mixture, side)
            bad = [self._entry(cache, 'bad', strand, k) for strand in (1, 0) for k in range(len(st['bad_part'][strand]))]
            MixtureFactory._add_side_annealings(mixture, bad)
        self._real = mixture
        return mixture


def load_mixture(path):
    """What ShelvedMixture(path) gives in the reference (:110-131): the saved mixture, read-only by convention."""
    return tmpStorage.from_shelf(path)


ShelvedMixture = load_mixture


class MixtureFactory(object):
    """PCR_Mixture.py:134-249: same constructor arguments, `_sort_annealings`, `create_PCR_mixture`."""

    def __init__(self, abort_event, primers, min_amplicon, max_amplicon, with_exonuclease=False,
                 include_side_annealings=False):
        self._abort_event = abort_event
        self._primers = primers
        self._min_amplicon = min_amplicon
        self._max_amplicon = max_amplicon
        self._with_exonuclease = with_exonuclease
        self._include_side_annealings = include_side_annealings

    def aborted(self):
        return self._abort_event is not None and self._abort_event.is_set()

    @staticmethod
    def _add_side_annealings(mixture, sorted_annealings):
        mixture.add_annealings(tuple((d, t) for _p, d, t in sorted_annealings))

    def _sort_annealings(self, hit, strand, annealings, good, bad):
        """Split every site's duplexes into those that can be extended (3' end annealed, or any when the
        polymerase has 3'-5' exonuclease activity) and those that cannot (:161-191).  `position` is the
        3' end of the primer: the site spans [pos-L+1, pos] on the forward strand, [pos, pos+L-1] on the reverse."""
        for pos, duplexes in annealings:
            length = duplexes[0].fwd_len
            if strand:
                template = Region(hit, pos - length + 1, pos, strand)
            else:
                template = Region(hit, pos, pos + length - 1, strand)
            good_here, bad_here = (pos, [], template), (pos, [], template)
            for duplex in duplexes:
                if not any(duplex.fwd_seq in primer for primer in self._primers):
                    continue                                  # no such primer in this system
                if self._with_exonuclease or duplex.have_3_matches:
                    good_here[1].append(duplex)
                else:
                    bad_here[1].append(duplex)
            if good_here[1]:
                good[strand].append(good_here)
            if bad_here[1]:
                bad.append(bad_here)

    def create_PCR_mixture(self, counter, hit, fwd_annealings, rev_annealings):
        """Pair forward and reverse sites into products (:194-249).  Returns a PCR_Mixture, or None when
        there is no product.

        The inner loop is kept as the reference has it, including what look like slips, because
        the products it reports depend on them:
          * `rev_start = rev_pi` stores an index into the SLICED list `good[0][rev_start:]`, so the
            window start can move backwards;
          * a pair shorter than min_amplicon `break`s the scan of reverse sites for this forward site
            just like a pair longer than max_amplicon does;
          * `added_positions[0]` remembers reverse sites by that same relative index.
        """
        if not fwd_annealings or not rev_annealings:
            counter.done()
            return None
        if isinstance(fwd_annealings, SitePool) and isinstance(rev_annealings, SitePool):
            return self._create_from_pools(counter, str(hit), fwd_annealings, rev_annealings)
        counter.set_work(3 if self._include_side_annealings else 2)
        hit = str(hit)
        mixture = PCR_Mixture(hit)
        good = ([], [])                  # [reverse, forward]
        bad = []
        self._sort_annealings(hit, 1, fwd_annealings, good, bad)
        self._sort_annealings(hit, 0, rev_annealings, good, bad)
        good[1].sort(key=lambda x: x[0])
        good[0].sort(key=lambda x: x[0])
        counter.count()
        products_added = False
        added = (set(), set())
        rev_start = 0
        for fwd_pi, (fwd_pos, fwd_dups, fwd_templ) in enumerate(good[1]):
            start = fwd_pos + 1
            for rev_pi, (rev_pos, rev_dups, rev_templ) in enumerate(good[0][rev_start:]):
                end = rev_pos - 1
                if start >= end:
                    rev_start = rev_pi
                    continue
                if not self._min_amplicon <= end - start + 1 <= self._max_amplicon:
                    break
                mixture.add_product(start, end, fwd_dups, rev_dups)
                if fwd_pi not in added[1]:
                    mixture.add_annealing(fwd_dups, fwd_templ)
                    added[1].add(fwd_pi)
                if rev_pi not in added[0]:
                    mixture.add_annealing(rev_dups, rev_templ)
                    added[0].add(rev_pi)
                products_added = True
        counter.count()
        if not products_added:
            return None
        if self._include_side_annealings:
            self._add_side_annealings(mixture, [a for i, a in enumerate(good[1]) if i not in added[1]])
            self._add_side_annealings(mixture, [a for i, a in enumerate(good[0]) if i not in added[0]])
            self._add_side_annealings(mixture, bad)
            counter.count()
        return mixture

    # ---- the same on device results: arrays in, device join, objects on demand ---------------------
    def _split_pool(self, pool):
        """_sort_annealings (:161-191) over a SitePool without touching a Duplex: per part the per-duplex
        flags (good = can be extended, known = the variant belongs to a primer of this system) and, per site,
        whether it has good / bad duplexes."""
        parts = []
        for part in pool.parts:
            site, variant, has3 = part.duplex_flags()
            table = part._table
            known_v = np.array([any(v in primer for primer in self._primers) for v in table.variants], dtype=bool)
            known = known_v[variant] if len(variant) else np.zeros(0, dtype=bool)
            good = np.ones(len(site), dtype=bool) if self._with_exonuclease else has3
            n = len(part)
            bounds = np.searchsorted(site, np.arange(n + 1))
            has_good = np.bincount(site[good & known], minlength=n) > 0
            has_bad = np.bincount(site[~good & known], minlength=n) > 0
            parts.append(dict(good=good, known=known, bounds=bounds, has_good=has_good, has_bad=has_bad,
                              pos=np.asarray(part.positions, dtype=np.int64)))
        return parts

    def _create_from_pools(self, counter, hit, fwd_pool, rev_pool):
        from . import TD_Functions as tdf
        counter.set_work(3 if self._include_side_annealings else 2)
        pools = {1: fwd_pool, 0: rev_pool}
        st = dict(pools=pools, include_side_annealings=self._include_side_annealings,
                  dup_good={}, dup_known={}, dup_bounds={}, good_part={}, good_site={}, good_pos={},
                  bad_part={}, bad_site={})
        for strand in (1, 0):
            parts = self._split_pool(pools[strand])
            st['dup_good'][strand] = [p['good'] for p in parts]
            st['dup_known'][strand] = [p['known'] for p in parts]
            st['dup_bounds'][strand] = [p['bounds'] for p in parts]
            for kind, flag in (('good', 'has_good'), ('bad', 'has_bad')):
                part_i = np.concatenate([np.full(int(p[flag].sum()), i, dtype=np.int64) for i, p in enumerate(parts)]
                                        or [np.zeros(0, dtype=np.int64)])
                site_i = np.concatenate([np.flatnonzero(p[flag]) for p in parts] or [np.zeros(0, dtype=np.int64)])
                pos = np.concatenate([p['pos'][p[flag]] for p in parts] or [np.zeros(0, dtype=np.int64)])
                if kind == 'good':                       # good[strand].sort(key=position): stable
                    order = np.argsort(pos, kind='stable')
                    part_i, site_i, pos = part_i[order], site_i[order], pos[order]
                    st['good_pos'][strand] = pos
                st[kind + '_part'][strand], st[kind + '_site'][strand] = part_i, site_i
        counter.count()
        pairs, rev_start = tdf._ctx().pair_products(st['good_pos'][1], st['good_pos'][0],
                                                    self._min_amplicon, self._max_amplicon)
        counter.count()
        if not len(pairs):
            return None
        st['pairs'], st['rev_start'] = pairs, rev_start
        if self._include_side_annealings:
            counter.count()
        return LazyMixture(hit, st)
