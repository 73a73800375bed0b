"""Dimer / Duplex result objects — host mirror of DegenPrimer/SecStructures.py:58-192,223-387.

The scan path never builds these in bulk: the device returns structure-of-arrays
dimer records (site, variant, match mask, dG) and `SiteDuplexes` materialises
`Duplex` / `Dimer` objects only when a consumer touches them (PCR_Mixture,
SinglePCR and the reports read `.dimers[i].K`, `.fwd_mismatch`, `.fwd_seq`, ...;
SURVEY §8b).  Constructing a `Duplex` by hand still works and evaluates its
dimers on the GPU through `TD_Functions.dimers_dG_corrected`.
"""
import numpy as np

from . import TD_Functions as tdf
from .SeqUtils import reverse_complement, rc_map   # noqa: F401  (rc_map re-exported like the reference)

max_dimer_dG = tdf.max_dimer_dG      # -3 kcal/mol, SecStructures.py:30
max_hairpin_dG = 3                   # SecStructures.py:31
min_K = tdf.min_K                    # 100, SecStructures.py:32


# bit positions of every byte value at every byte offset of a 64-bit mask, and every byte value with a
# zero bit interleaved between its bits (bit b -> bit 2b): table-driven forms of the two inner loops below
_BYTE_BITS = [[tuple(8 * k + b for b in range(8) if v >> b & 1) for v in range(256)] for k in range(8)]
_BYTE_SPREAD = [sum(1 << (2 * b) for b in range(8) if v >> b & 1) for v in range(256)]


def mask_to_matches(mask):
    """Sorted tuple of the set bit positions of a match mask (<= 64 bits)."""
    mask = int(mask)
    out = ()
    k = 0
    while mask:
        out += _BYTE_BITS[k][mask & 255]
        mask >>= 8
        k += 1
    return out


def _structure_key_of_mask(mask, offset):
    """`structure_key` as ONE integer: bit (2f - offset + 64) set for every match f.  Equal integers <=> equal
    key tuples (both list the same increasing values), so it de-duplicates exactly like the tuple does."""
    mask = int(mask)
    spread = 0
    shift = 0
    while mask:
        spread |= _BYTE_SPREAD[mask & 255] << shift
        mask >>= 8
        shift += 16
    return spread << (64 - offset)


class Dimer(object):
    """Match-index structure of a dimer without bulges (SecStructures.py:58-192)."""
    __slots__ = ['fwd_matches', 'offset', 'num_matches', 'conversion_degree', 'dG', 'K', 'fwd_mismatch']

    def __init__(self, fwd_matches=None, offset=0):
        if fwd_matches:
            self.fwd_matches = tuple(sorted(fwd_matches))
            self.offset = offset
            self.num_matches = len(self.fwd_matches)
        else:
            self.fwd_matches = tuple()
            self.offset = None
            self.num_matches = 0
        self.conversion_degree = None
        self.dG = None
        self.K = None
        self.fwd_mismatch = None

    def __hash__(self):
        return hash((self.fwd_matches, self.offset))

    def __eq__(self, other):
        return hash(self) == hash(other)

    def __ne__(self, other):
        return hash(self) != hash(other)

    def __bool__(self):
        return self.num_matches > 0

    def __getstate__(self):
        return tuple(getattr(self, s) for s in self.__slots__)

    def __setstate__(self, state):
        for s, v in zip(self.__slots__, state):
            setattr(self, s, v)

    rev_matches = property(lambda self: np.array(self.fwd_matches) - self.offset)
    fwd_min = property(lambda self: self.fwd_matches[0])
    rev_min = property(lambda self: self.fwd_matches[0] - self.offset)
    fwd_max = property(lambda self: self.fwd_matches[-1])
    rev_max = property(lambda self: self.fwd_matches[-1] - self.offset)

    @property
    def alternatives(self):
        """Every contiguous range of match runs except the full one, in the
        reference's order (SecStructures.py:126-152)."""
        if self.num_matches < 2:
            return []
        runs = [[self.fwd_matches[0]]]
        for a, b in zip(self.fwd_matches, self.fwd_matches[1:]):
            if b - a > 1:
                runs.append([])
            runs[-1].append(b)
        last = len(runs) - 1
        out = []
        for l in range(last):
            for r in range(last, l, -1):
                out.append(Dimer([i for run in runs[l:r] for i in run], self.offset))
            out.append(Dimer([i for run in runs[l + 1:] for i in run], self.offset))
        return out

    def add(self, fwd, rev):
        if self.offset is None:
            self.offset = fwd - rev
        elif self.offset != fwd - rev:
            raise ValueError('Dimer.add: fwd and rev indexes should be at the same distance as all '
                             'other match pairs. Asymmetrical loops are not allowed.')
        self.fwd_matches = tuple(sorted(self.fwd_matches + (fwd,)))
        self.num_matches += 1

    @classmethod
    def from_sequences(cls, forward, revcomp, offset=0):
        """Left-aligned (or offset) character comparison (SecStructures.py:173-191)."""
        d = cls()
        d.offset = offset
        if offset > 0:
            rng = range(offset, min(len(forward), len(revcomp) + offset))
        else:
            rng = range(0, min(len(forward) - offset, len(revcomp) + offset))
        d.fwd_matches = tuple(i for i in rng if forward[i] == revcomp[i - offset])
        d.num_matches = len(d.fwd_matches)
        return d


class Duplex(object):
    """A DNA:DNA duplex and its stable dimers (SecStructures.py:223-387)."""
    __slots__ = ['_fwd_sequence', '_rev_sequence', 'fwd_len', 'rev_len', 'name', '_dimers',
                 '_nonzero', '_fwd_matches', '_fwd_mismatches', 'fwd_3_overhang']

    def __init__(self, fwd_seq, rev_seq, name='', dimer=None, revcomp=False):
        self.name = name
        self.fwd_len = len(fwd_seq)
        self.rev_len = len(rev_seq)
        self._fwd_sequence = fwd_seq
        self._rev_sequence = reverse_complement(rev_seq) if revcomp else rev_seq
        if not dimer:
            dimer = Dimer.from_sequences(fwd_seq, rev_seq if revcomp else reverse_complement(rev_seq))
        self._dimers = [dimer] + dimer.alternatives
        self._fwd_matches = []
        self._fwd_mismatches = []
        self.fwd_3_overhang = None
        self._nonzero = False
        self._recalculate()

    @classmethod
    def _from_device(cls, name, fwd_seq, window, masks, dGs, T):
        """Build from the rows dp_duplex_dg returned for one (site, variant):
        already filtered (dG <= -3, K >= 100) and stably sorted by dG."""
        self = cls.__new__(cls)
        self.name = name
        self.fwd_len = self.rev_len = len(fwd_seq)
        self._fwd_sequence = fwd_seq
        self._rev_sequence = reverse_complement(window)
        dimers = []
        for mask, dG in zip(masks, dGs):
            d = Dimer(mask_to_matches(mask), 0)
            d.dG = float(dG)
            d.K = tdf.equilibrium_constant(d.dG, T)
            d.fwd_mismatch = d.fwd_max < self.fwd_len - 1
            dimers.append(d)
        self._dimers = tuple(dimers)
        self._nonzero = bool(dimers)
        self._fwd_matches = []
        self._fwd_mismatches = []
        self.fwd_3_overhang = None
        self._index_3_ends()
        return self

    def _recalculate(self):
        """SecStructures.py:294-324; dG of all candidate dimers in one launch."""
        cand = [d for d in self._dimers if d]
        kept = []
        if cand:
            dGs = tdf.dimers_dG_corrected(cand, [self._fwd_sequence] * len(cand),
                                          [self._rev_sequence] * len(cand))
            for d, g in zip(cand, dGs):
                d.dG = float(g)
                if d.dG > max_dimer_dG:
                    continue
                d.K = tdf.equilibrium_constant(d.dG, tdf.PCR_P.PCR_T)
                if d.K < min_K:
                    continue
                d.fwd_mismatch = d.fwd_max < self.fwd_len - 1
                kept.append(d)
        self._nonzero = bool(kept)
        if not self._nonzero:
            self._dimers = tuple()
            return
        self._dimers = tuple(sorted(kept, key=lambda x: x.dG))
        self._index_3_ends()

    def _index_3_ends(self):
        if not self._dimers:
            return
        fwd_tail = self.fwd_len - 1 - self._dimers[0].fwd_max
        rev_tail = self.rev_len - 1 - self._dimers[0].rev_max
        self.fwd_3_overhang = fwd_tail - rev_tail if fwd_tail > rev_tail else 0
        for i, d in enumerate(self._dimers):
            (self._fwd_mismatches if d.fwd_max < self.fwd_len - 1 else self._fwd_matches).append(i)

    def __bool__(self):
        return self._nonzero

    def __hash__(self):
        return hash((self._fwd_sequence, self._rev_sequence, self._dimers[0]))

    def __eq__(self, other):
        return hash(self) == hash(other)

    def __ne__(self, other):
        return hash(self) != hash(other)

    def __getstate__(self):
        return tuple(getattr(self, s) for s in self.__slots__)

    def __setstate__(self, state):
        for s, v in zip(self.__slots__, state):
            setattr(self, s, v)

    fwd_seq = property(lambda self: self._fwd_sequence)
    rev_seq = property(lambda self: self._rev_sequence)
    dimer = property(lambda self: self._dimers[0])
    dimers = property(lambda self: self._dimers)
    dG = property(lambda self: self._dimers[0].dG)
    K = property(lambda self: self._dimers[0].K)
    fwd_3_matches = property(lambda self: [self._dimers[i] for i in self._fwd_matches])
    fwd_3_mismatches = property(lambda self: [self._dimers[i] for i in self._fwd_mismatches])
    have_3_matches = property(lambda self: bool(self._fwd_matches))

    def __str__(self):
        s = '%d variants of annealing sorted by stability:\n' % len(self._dimers)
        for i, d in enumerate(self._dimers):
            s += '\n%d:\n' % (i + 1)
            s += self.print_dimer(self._fwd_sequence, self._rev_sequence, d)
            s += '\n'
        return s + '\n'

    __repr__ = __str__

    def print_most_stable(self, include_fwd_3_mismatch=True):
        s = ''
        if not self._dimers:
            return s
        if not include_fwd_3_mismatch:
            if self._fwd_matches:
                return self.print_dimer(self._fwd_sequence, self._rev_sequence,
                                        self._dimers[self._fwd_matches[0]])
            s += "There's no annealing variant without 3' mismatches\n"
        return s + self.print_dimer(self._fwd_sequence, self._rev_sequence, self._dimers[0])

    @classmethod
    def print_dimer(cls, fwd_sequence, rev_sequence, dimer):
        """ASCII rendering, SecStructures.py:355-386."""
        if not dimer:
            return fwd_sequence + '\n' + rev_sequence[::-1] + '\n'
        bars = [' '] * len(fwd_sequence)
        for m in dimer.fwd_matches:
            bars[m] = '|'
        bars = ''.join(bars)
        s = ''
        if dimer.offset > 0:
            s += "5' " + fwd_sequence + " 3'\n   " + bars + '\n'
            s += ' ' * dimer.offset + "3' " + rev_sequence[::-1] + " 5'\n"
        else:
            pad = ' ' * abs(dimer.offset) if dimer.offset < 0 else ''
            s += pad + "5' " + fwd_sequence + " 3'\n   " + pad + bars + '\n'
            s += "3' " + rev_sequence[::-1] + " 5'\n"
        if dimer.dG is not None:
            s += 'dG(%.1fC) = %.2f kcal/mol\n' % (tdf.PCR_P.PCR_T, dimer.dG)
        if dimer.conversion_degree is not None:
            s += 'conversion degree = %.4f%%\n' % (dimer.conversion_degree * 100)
        return s


class DuplexTable(object):
    """Structure-of-arrays result of one dp_duplex_dg call, shared by all sites."""

    def __init__(self, dimers, windows, variants, names, T):
        self.dimers = dimers        # DIMER_DTYPE rows grouped by (hit, variant)
        self.windows = windows      # list: hit index -> window string (strand orientation)
        self.variants = variants    # variant sequences of the primer
        self.names = names          # variant ids
        self.T = T                  # PCR_T the rows were computed at


class SiteDuplexes(object):
    """Tuple-like view of the duplexes at one annealing site (the second item of
    the `(position, duplexes)` pairs of DegenPrimer/SearchEngine.py:106-118)."""
    __slots__ = ['_table', '_hit', '_lo', '_hi', '_cache']

    def __init__(self, table, hit, lo, hi):
        self._table, self._hit, self._lo, self._hi = table, hit, lo, hi
        self._cache = None

    def _materialise(self):
        if self._cache is None:
            t = self._table
            rows = t.dimers[self._lo:self._hi]
            out = []
            cut = np.flatnonzero(np.diff(rows['variant'])) + 1
            for seg in np.split(np.arange(len(rows)), cut):
                v = int(rows['variant'][seg[0]])
                out.append(Duplex._from_device(t.names[v], t.variants[v], t.windows[self._hit],
                                               rows['match_mask'][seg], rows['dG'][seg], t.T))
            self._cache = tuple(out)
        return self._cache

    def __len__(self):
        rows = self._table.dimers['variant'][self._lo:self._hi]
        return int(np.count_nonzero(np.diff(rows)) + 1) if len(rows) else 0

    def __bool__(self):
        return self._hi > self._lo

    def __iter__(self):
        return iter(self._materialise())

    def __getitem__(self, i):
        return self._materialise()[i]

    def __eq__(self, other):
        return tuple(self) == tuple(other)

    def __hash__(self):
        return hash(tuple(self))

    def __add__(self, other):
        return tuple(self) + tuple(other)

    def __repr__(self):
        return repr(tuple(self))

    def __reduce__(self):
        return (tuple, (list(self._materialise()),))


class LazyWindows(object):
    """hit index -> window string (strand orientation), fetched from HBM in ONE dp_hit_windows
    call on first use and kept for every table that shares the hits."""

    def __init__(self, tmpl, hits, primer_lens):
        self._tmpl, self._hits, self._lens = tmpl, hits, list(primer_lens)
        self._w = None

    def _fetch(self):
        if self._w is None:
            self._w = self._tmpl.ctx.hit_windows(self._tmpl, self._hits, self._lens)
            self._tmpl = None
        return self._w

    def __getitem__(self, h):
        return self._fetch()[h]

    def __len__(self):
        return len(self._hits)


class SitePool(object):
    """The annealing sites of several primers on one strand, pooled (PCR_ProductsFinder._combine_annealings,
    DegenPrimer/PCR_ProductsFinder.py:57-65) but kept as the SiteLists they came in: list-like for every
    consumer of `[(position, duplexes), ...]`, structure-of-arrays for MixtureFactory's device join."""

    def __init__(self, parts=()):
        self.parts = [p for p in parts if len(p)]

    def extend(self, part):
        if len(part):
            self.parts.append(part)

    def __len__(self):
        return sum(len(p) for p in self.parts)

    def __bool__(self):
        return bool(self.parts)

    def __iter__(self):
        for p in self.parts:
            for item in p:
                yield item

    def __getitem__(self, i):
        if isinstance(i, slice):
            return list(self)[i]
        if i < 0:
            i += len(self)
        for p in self.parts:
            if i < len(p):
                return p[i]
            i -= len(p)
        raise IndexError('SitePool index out of range')

    def __eq__(self, other):
        return list(self) == list(other)

    __hash__ = None

    def __reduce__(self):
        return (SitePool, (self.parts,))


def _rebuild_site_list(positions, counts, rows, windows, variants, names, T):
    table = DuplexTable(rows, windows, variants, names, T)
    hi = np.cumsum(counts)
    return SiteList(table, np.arange(len(positions)), positions, hi - counts, hi)


class SiteList(object):
    """List-like `[(position, duplexes), ...]` of one primer on one strand of one template — the
    fwd_results / rev_results of DegenPrimer/SearchEngine.py:120-157 — kept as arrays: the
    (position, SiteDuplexes) pairs are made when the list is iterated or indexed, Duplex objects only
    when a SiteDuplexes is looked into.  Pickles as arrays (what `find` puts on the shelf)."""
    __slots__ = ['_table', '_hits', '_pos', '_lo', '_hi']

    def __init__(self, table, hits, positions, lo, hi):
        self._table = table
        self._hits = np.asarray(hits)
        self._pos = np.asarray(positions)
        self._lo, self._hi = np.asarray(lo), np.asarray(hi)

    def __len__(self):
        return len(self._pos)

    def __bool__(self):
        return len(self._pos) > 0

    def _item(self, i):
        return (int(self._pos[i]), SiteDuplexes(self._table, int(self._hits[i]), int(self._lo[i]), int(self._hi[i])))

    def __iter__(self):
        for i in range(len(self._pos)):
            yield self._item(i)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return SiteList(self._table, self._hits[i], self._pos[i], self._lo[i], self._hi[i])
        if i < 0:
            i += len(self._pos)
        if not 0 <= i < len(self._pos):
            raise IndexError('SiteList index out of range')
        return self._item(i)

    @property
    def positions(self):
        return self._pos

    def sort(self, key=None, reverse=False):
        """Order by position (the only order the callers ask for, SearchEngine.py:153-154)."""
        order = np.argsort(self._pos, kind='stable')
        if reverse:
            order = order[::-1]
        self._hits, self._pos = self._hits[order], self._pos[order]
        self._lo, self._hi = self._lo[order], self._hi[order]

    def __eq__(self, other):
        return list(self) == list(other)

    def __ne__(self, other):
        return not self == other

    __hash__ = None

    def __add__(self, other):
        return list(self) + list(other)

    def __repr__(self):
        return 'SiteList(%d sites)' % len(self)

    def duplex_flags(self):
        """Per duplex ((site, variant) group of dimer rows) of this list, as arrays and without building any
        object: (site index in this list, variant index, has a dimer whose 3' end is annealed).  The last one is
        Duplex.have_3_matches (SecStructures.py:326-339: some dimer with fwd_max == fwd_len - 1)."""
        t = self._table
        n = (self._hi - self._lo).astype(np.int64)
        if not len(n) or not int(n.sum()):
            z = np.zeros(0, dtype=np.int64)
            return z, z, np.zeros(0, dtype=bool)
        site = np.repeat(np.arange(len(n)), n)
        first = np.concatenate(([0], np.cumsum(n)[:-1]))
        rows = np.repeat(self._lo.astype(np.int64) - first, n) + np.arange(int(n.sum()))
        variant = t.dimers['variant'][rows].astype(np.int64)
        p_len = len(t.variants[0])
        bit = ((t.dimers['match_mask'][rows] >> np.uint64(p_len - 1)) & np.uint64(1)).astype(bool)
        new = np.ones(len(rows), dtype=bool)
        new[1:] = (site[1:] != site[:-1]) | (variant[1:] != variant[:-1])
        starts = np.flatnonzero(new)
        return site[starts], variant[starts], np.logical_or.reduceat(bit, starts)

    def __reduce__(self):
        t = self._table
        counts = (self._hi - self._lo).astype(np.int64)
        rows = (np.concatenate([t.dimers[a:b] for a, b in zip(self._lo, self._hi)])
                if len(counts) else t.dimers[:0].copy())
        windows = [t.windows[int(h)] for h in self._hits]
        return (_rebuild_site_list, (np.array(self._pos), counts, rows, windows, t.variants, t.names, t.T))


# ---------------------------------------------------------------------------------------------
# self-dimers, cross-dimers and hairpins of primer variants
# (DegenPrimer/SecStructures.py:389-716).  Two kernels: dp_find_dimers (one warp per sequence pair and
# relative position) and dp_find_hairpins (one warp per hairpin seed).
class Hairpin(object):
    """Stem of a hairpin (SecStructures.py:194-220): fwd_matches ascending, rev_matches descending,
    fwd_matches[i] pairs with rev_matches[i]."""
    __slots__ = ['fwd_matches', 'rev_matches', 'conversion_degree', 'dG', 'K']

    def __init__(self, fwd_matches, rev_matches):
        if len(fwd_matches) != len(rev_matches):
            raise ValueError('SecStructures.Dimer: number of forward matches should equal that of reverse matches.')
        self.fwd_matches = tuple(sorted(fwd_matches))
        self.rev_matches = tuple(sorted(rev_matches, reverse=True))
        self.conversion_degree = None
        self.dG = None
        self.K = None

    def __getstate__(self):
        return tuple(getattr(self, s) for s in self.__slots__)

    def __setstate__(self, state):
        for s, v in zip(self.__slots__, state):
            setattr(self, s, v)


class Reaction(object):
    """The four fields of Equilibrium.Reaction (DegenPrimer/Equilibrium.py:33-62) that
    SecStructures.compose_reactions fills; Equilibrium itself is out of scope."""
    __slots__ = ['constant', 'Ai', 'Bi', 'type']

    def __init__(self, K, Ai, Bi, r_type):
        self.constant, self.Ai, self.Bi, self.type = K, Ai, Bi, r_type

    def __str__(self):
        return '[K=%f; Ai: %s; Bi: %s; type: %s]' % (self.constant, str(self.Ai), str(self.Bi), self.type)


class DimerSet(object):
    """What SecStructures._find_dimers leaves behind for ONE sequence pair
    (DegenPrimer/SecStructures.py:609-668): `dimers` — the most stable structure of every
    relative position with dG < max_dimer_dG, de-duplicated by the reference's structure key,
    in the reference's scan order; `fronts` — for self-dimers, the full (un-trimmed) dimer of every
    position that has at least one match, first of each structure key (the reference's hairpin
    seeds, :643-645); the 3'-dimer counter and the most stable 3' dimer (:660-667)."""
    __slots__ = ['seq1', 'seq2', 'dimers', 'fronts', 'n_3prim_dimers', 'min_3prim_dimer', 'min_3prim_dimer_dG']

    def __init__(self, seq1, seq2):
        self.seq1, self.seq2 = seq1, seq2
        self.dimers, self.fronts = [], []
        self.n_3prim_dimers = 0
        self.min_3prim_dimer = None
        self.min_3prim_dimer_dG = None

    def __bool__(self):
        return bool(self.dimers)

    def __len__(self):
        return len(self.dimers)


class HairpinSet(object):
    """What SecStructures._find_hairpins leaves behind for one sequence (:676-716)."""
    __slots__ = ['seq', 'hairpins', 'n_3prim_hairpins', 'min_3prim_hairpin', 'min_3prim_hairpin_dG']

    def __init__(self, seq):
        self.seq = seq
        self.hairpins = []
        self.n_3prim_hairpins = 0
        self.min_3prim_hairpin = None
        self.min_3prim_hairpin_dG = None

    def __bool__(self):
        return bool(self.hairpins)

    def __len__(self):
        return len(self.hairpins)


def structure_key(dimer):
    """The tuple inside SecStructures._structure_hash (:597-605): `fwd_matches + rev_matches` adds a
    tuple to a numpy array, i.e. ELEMENT-WISE — the key is (2*f - offset for f in fwd_matches)."""
    return tuple(2 * f - dimer.offset for f in dimer.fwd_matches)


def find_dimers_batch(seqs, pairs, return_ms=False):
    """SecStructures._find_dimers for many sequence pairs in ONE kernel launch (dp_find_dimers:
    one warp per (pair, relative position); [dimer] + alternatives evaluated by the lanes).

    seqs: list of str / Seq (5'->3', <= 64 nt); pairs: iterable of (i, j) indices into seqs
    (i == j or equal sequences: self-dimers).  Returns a list of DimerSet, one per pair."""
    strs = [str(s).upper() for s in seqs]
    pairs = np.asarray(list(pairs), dtype=np.uint32).reshape(-1, 2)
    rec, ms = tdf._ctx().find_dimers(strs, pairs, tdf.conditions())
    out = _dimer_sets_from_records(strs, pairs, rec)
    return (out, ms) if return_ms else out


def _dimer_sets_from_records(strs, pairs, rec):
    """Host half of `find_dimers_batch`: the dense [pair][front] records of dp_find_dimers -> DimerSet objects
    (de-duplication by the reference's structure key, hairpin seeds of self pairs, 3'-dimer bookkeeping)."""
    if np.isnan(rec['best_dG']).any():
        raise ValueError('UnifiedNN.int_dPar_37: sequence is not in the database')
    la = np.array([len(s) for s in strs], dtype=np.int64)
    plist = pairs.tolist()                                   # plain ints: numpy scalars are slow list indices
    out = [DimerSet(strs[i], strs[j]) for i, j in plist]
    is_self = np.array([strs[i] == strs[j] for i, j in plist], dtype=bool) if plist else np.zeros(0, bool)
    # only the entries that lead to an object are visited: stable structures, and the seeds of self pairs
    want = (rec['best_mask'] != 0) | ((rec['full_mask'] != 0) & is_self[:, None])
    pp, ff = np.nonzero(want)                                # row-major: pair by pair, fronts ascending
    fwd_len = la[pairs[pp, 0]] if len(pp) else np.zeros(0, np.int64)
    offsets = np.maximum(0, fwd_len - ff - 1) - np.maximum(0, ff - fwd_len + 1)
    full, best, dGs = rec['full_mask'][pp, ff], rec['best_mask'][pp, ff], rec['best_dG'][pp, ff]
    self_list = is_self.tolist()
    seen, seen_fronts, last = None, None, -1
    for p, offset, fm, bm, dG, flen in zip(pp.tolist(), offsets.tolist(), full.tolist(), best.tolist(),
                                           dGs.tolist(), fwd_len.tolist()):
        if p != last:
            seen, seen_fronts, last = set(), set(), p
        ds = out[p]
        if self_list[p]:
            key = _structure_key_of_mask(fm, offset)
            if key not in seen_fronts:                                        # SecStructures.py:643-645
                seen_fronts.add(key)
                ds.fronts.append(Dimer(mask_to_matches(fm), offset))
        if not bm:
            continue
        key = _structure_key_of_mask(bm, offset)
        if key in seen:
            continue
        seen.add(key)
        d = Dimer(mask_to_matches(bm), offset)
        d.dG = dG
        ds.dimers.append(d)
        if d.fwd_matches[-1] > flen - 4 or d.rev_min < 3:                     # SecStructures.py:660-667
            ds.n_3prim_dimers += 1
            if ds.min_3prim_dimer_dG is None or ds.min_3prim_dimer_dG > d.dG:
                ds.min_3prim_dimer_dG = d.dG
                ds.min_3prim_dimer = d
    return out


def find_dimers(seq1, seq2=None):
    """One pair: cross-dimers of seq1 and seq2, or self-dimers of seq1 (SecStructures.py:576-588)."""
    if seq2 is None or str(seq1) == str(seq2):
        return find_dimers_batch([seq1], [(0, 0)])[0]
    return find_dimers_batch([seq1, seq2], [(0, 1)])[0]


def trim_hairpin_seed(fwd_matches):
    """The while loop of SecStructures._find_hairpins (:684-690) on the sorted match list m: after t
    rounds the 5' arm is m[:k-t] and the 3' arm m[t:]; pairs are dropped from the middle until the arms
    neither cross nor come closer than 4 (3 unpaired bases).  Returns (fwd arm, rev arm)."""
    m = sorted(fwd_matches)
    k, t = len(m), 0
    while t < k and (m[t] < m[k - 1 - t] or m[t] - m[k - 1 - t] < 4):
        t += 1
    return m[:k - t], m[t:]


def _hairpin_seeds(dimer_sets):
    """(seed_seq, seed_mask, seed_offset) for dp_find_hairpins: every hairpin seed of every self-dimer set,
    trimmed (SecStructures.py:684-690,695)."""
    seed_seq, seed_mask, seed_off = [], [], []
    for si, ds in enumerate(dimer_sets):
        L = len(ds.seq1)
        for seed in ds.fronts:
            fwd, rev = trim_hairpin_seed(seed.fwd_matches)
            if not fwd:
                continue
            m = 0
            for f in fwd:
                m |= 1 << f
            seed_seq.append(si)
            seed_mask.append(m)
            seed_off.append(fwd[0] - (L - 1 - rev[-1]))
    return seed_seq, seed_mask, seed_off


def _hairpin_sets_from_records(seqs, seed_seq, seed_off, rec):
    """Host half of `find_hairpins_batch`: the per-seed records of dp_find_hairpins -> HairpinSet objects
    (de-duplication by structure key, 3'-hairpin bookkeeping, SecStructures.py:700-714)."""
    out = [HairpinSet(s) for s in seqs]
    if not len(seed_seq):
        return out
    if np.isnan(rec['best_dG']).any():
        raise ValueError('UnifiedNN.int_dPar_37: sequence is not in the database')
    seen, last = None, -1
    for si, off, bm, dG in zip(seed_seq, seed_off, rec['best_mask'].tolist(), rec['best_dG'].tolist()):
        if si != last:
            seen, last = set(), si
        if not bm:
            continue
        hs = out[si]
        L = len(hs.seq)
        fm = mask_to_matches(bm)
        h = Hairpin(fm, [L - 1 - (f - off) for f in fm])
        h.dG = dG
        key = h.fwd_matches + h.rev_matches                               # _structure_hash: concatenation
        if key in seen:
            continue
        seen.add(key)
        hs.hairpins.append(h)
        if h.rev_matches[0] > L - 4:
            hs.n_3prim_hairpins += 1
            if hs.min_3prim_hairpin_dG is None or hs.min_3prim_hairpin_dG > dG:
                hs.min_3prim_hairpin_dG = dG
                hs.min_3prim_hairpin = h
    return out


def find_hairpins_batch(dimer_sets, return_ms=False):
    """SecStructures._find_hairpins for the self-dimer seeds (`DimerSet.fronts`) of many sequences in
    ONE kernel launch (dp_find_hairpins: one warp per seed).  Returns a list of HairpinSet."""
    seqs = [ds.seq1 for ds in dimer_sets]
    seed_seq, seed_mask, seed_off = _hairpin_seeds(dimer_sets)
    rec, ms = None, 0.0
    if seed_seq:
        rec, ms = tdf._ctx_with_hairpin_tables().find_hairpins(seqs, seed_seq, seed_mask, seed_off, tdf.conditions(),
                                                               max_hairpin_dG)
    out = _hairpin_sets_from_records(seqs, seed_seq, seed_off, rec)
    return (out, ms) if return_ms else out


class SecStructures(object):
    """Self-dimers and hairpins of one sequence, or cross-dimers of two
    (DegenPrimer/SecStructures.py:389-871): same constructor, same attributes, query methods and report
    text (`str()`, `formatFull()`, `formatShort()`).
    `SecStructures.batch(records, pairs)` builds many of them from two kernel launches."""

    def __init__(self, seq_rec1, seq_rec2=None, _dimer_set=None, _hairpin_set=None):
        self._seq_rec1 = seq_rec1
        self._seq1 = seq_rec1.seq
        if seq_rec2 and str(seq_rec1.seq) != str(seq_rec2.seq):
            self._seq_rec2 = seq_rec2
            self._seq2 = seq_rec2.seq
        else:
            self._seq_rec2 = None
            self._seq2 = None
        if _dimer_set is None:
            _dimer_set = find_dimers(str(self._seq1), None if self._seq2 is None else str(self._seq2))
        if self._seq2 is None and _hairpin_set is None:
            _hairpin_set = find_hairpins_batch([_dimer_set])[0]
        self._fill(_dimer_set, _hairpin_set)

    def _fill(self, ds, hs):
        dimers = dict((self._structure_hash(d), d) for d in ds.dimers)
        self._3prim_dimers = ds.n_3prim_dimers
        self._min_3prim_dimer_dG = ds.min_3prim_dimer_dG
        self._min_3prim_dimer = ds.min_3prim_dimer
        self._3prim_hairpins = 0
        self._min_3prim_hairpin_dG = None
        self._min_3prim_hairpin = None
        if self._seq2 is not None:
            self._seq1_dimers = None
            self._seq1_hairpins = None
            self._cross_dimers = dimers
        else:
            self._seq1_dimers = dimers
            self._seq1_hairpins = dict((self._structure_hash(h), h) for h in hs.hairpins)
            self._cross_dimers = None
            self._3prim_hairpins = hs.n_3prim_hairpins
            self._min_3prim_hairpin_dG = hs.min_3prim_hairpin_dG
            self._min_3prim_hairpin = hs.min_3prim_hairpin

    @classmethod
    def batch(cls, records, pairs):
        """[SecStructures(records[i], records[j]) for i, j in pairs] (i == j: self structures) with one
        dp_find_dimers and one dp_find_hairpins launch for all of them (AllSecStructures.py:90-121)."""
        pairs = [(int(i), int(j)) for i, j in pairs]
        sets = find_dimers_batch([str(r.seq) for r in records], pairs)
        selfs = [k for k, (i, j) in enumerate(pairs) if str(records[i].seq) == str(records[j].seq)]
        hsets = dict(zip(selfs, find_hairpins_batch([sets[k] for k in selfs])))
        return [cls(records[i], records[j], _dimer_set=sets[k], _hairpin_set=hsets.get(k))
                for k, (i, j) in enumerate(pairs)]

    def _structure_hash(self, struct):
        rev = struct.rev_matches
        key = struct.fwd_matches + rev if isinstance(rev, tuple) else structure_key(struct)
        if not self._seq_rec2:
            return hash((str(self._seq1), tuple(key)))
        return hash((str(self._seq1), str(self._seq2), tuple(key)))

    def __bool__(self):
        if self._seq2 and self._cross_dimers:
            return True
        return bool(self._seq1_dimers or self._seq1_hairpins)

    def __contains__(self, s_hash):
        if not self._seq_rec2:
            return s_hash in self._seq1_dimers or s_hash in self._seq1_hairpins
        return s_hash in self._cross_dimers

    def compose_reactions(self):
        """SecStructures.py:439-466."""
        reactions = dict()
        T = tdf.PCR_P.PCR_T
        if not self._seq_rec2:
            for h, D in (self._seq1_dimers or {}).items():
                reactions[h] = Reaction(tdf.equilibrium_constant(D.dG, T), str(self._seq1), None, '2A')
            for h, H in (self._seq1_hairpins or {}).items():
                reactions[h] = Reaction(tdf.equilibrium_constant(H.dG, T), str(self._seq1), None, 'A')
        else:
            for h, D in (self._cross_dimers or {}).items():
                reactions[h] = Reaction(tdf.equilibrium_constant(D.dG, T), str(self._seq1), str(self._seq2), 'AB')
        return reactions

    def set_conversion_degree(self, s_hash, r):
        """SecStructures.py:469-482."""
        if not self._seq_rec2:
            for table in (self._seq1_dimers, self._seq1_hairpins):
                if s_hash in table:
                    table[s_hash].conversion_degree = r
                    return True
        elif s_hash in self._cross_dimers:
            self._cross_dimers[s_hash].conversion_degree = r
            return True

    def max_conversion_degree(self):
        if self._seq_rec2 and self._cross_dimers:
            return max(s.conversion_degree for s in self._cross_dimers.values())
        CG = 0
        if self._seq1_dimers:
            CG += max(s.conversion_degree for s in self._seq1_dimers.values())
        if self._seq1_hairpins:
            CG += max(s.conversion_degree for s in self._seq1_hairpins.values())
        return CG if CG else None

    def dimerMin_dG(self):
        if self._seq_rec2 and self._cross_dimers:
            return min(s.dG for s in self._cross_dimers.values())
        if self._seq1_dimers:
            return min(s.dG for s in self._seq1_dimers.values())
        return None

    def dimerMin_3prim_dG(self):
        return self._min_3prim_dimer_dG if self._min_3prim_dimer_dG else None

    def hairpinMin_dG(self):
        if self._seq_rec2 or not self._seq1_hairpins:
            return None
        return min(s.dG for s in self._seq1_hairpins.values())

    def hairpinMin_3prim_dG(self):
        return self._min_3prim_hairpin_dG if self._min_3prim_hairpin_dG else None

    # ---- report text (SecStructures.py:528-567,718-871) -----------------------------------------
    def formatFull(self):
        return self._format_str(full=True)

    def formatShort(self):
        return self._format_str(full=False)

    def __str__(self):
        return self._format_str()

    def _format_str(self, full=True):
        """All structures (`full`) or counts, minimum dG and the most stable ones; cross-dimers for two
        sequences, self-dimers and hairpins for one."""
        from .Product import hr
        show_dimers = self._format_dimers if full else self._format_dimers_short
        show_hairpins = self._format_hairpins if full else self._format_hairpins_short
        by_dG = lambda table: sorted(table.values(), key=lambda x: x.dG)
        seq1 = str(self._seq1)
        out = ['\n']
        if self._seq2:
            out.append(hr(' %s vs %s cross-dimers ' % (self._seq_rec1.id, self._seq_rec2.id), symbol='='))
            out.append(show_dimers(by_dG(self._cross_dimers), seq1, str(self._seq2)))
        else:
            out.append(hr(' %s: %s ' % (self._seq_rec1.id, seq1), symbol='='))
            out.append(hr(' self-dimers '))
            out.append(show_dimers(by_dG(self._seq1_dimers), seq1, seq1))
            out.append(hr(' hairpins '))
            out.append(show_hairpins(by_dG(self._seq1_hairpins), seq1))
        out.append(hr('', symbol='='))
        return ''.join(out)

    @staticmethod
    def _is_3prim_dimer(dimer, seq1):
        return dimer.fwd_matches[-1] > len(seq1) - 4 or dimer.rev_min < 3

    @classmethod
    def _format_dimer(cls, dimer, seq1, seq2):
        text = Duplex.print_dimer(str(seq1), str(seq2), dimer)
        if cls._is_3prim_dimer(dimer, seq1):
            text += "(3'-dimer)\n"
        return text + '\n'

    def _format_dimers_header(self, dimers):
        if not dimers:
            return 'No dimers found.\n\n'
        lines = ['Dimers count: %d' % len(dimers)]
        if self._3prim_dimers:
            lines.append("3'-dimers count: %d" % self._3prim_dimers)
        lines.append('Most stable dimer: %.2f kcal/mol' % self.dimerMin_dG())
        if self._3prim_dimers:
            lines.append("Most stable 3' dimer: %.2f kcal/mol" % self.dimerMin_3prim_dG())
        return '\n'.join(lines) + '\n\n'

    def _format_dimers(self, dimers, seq1, seq2):
        if not dimers:
            return 'No dimers found.\n\n'
        return self._format_dimers_header(dimers) + ''.join(self._format_dimer(d, seq1, seq2) for d in dimers)

    def _format_dimers_short(self, dimers, seq1, seq2):
        """Header, the most stable dimer and the most stable 3' dimer (once, if they are the same)."""
        if not dimers:
            return 'No dimers found.\n\n'
        text = self._format_dimers_header(dimers)
        if self.dimerMin_dG() and dimers[0] != self._min_3prim_dimer:
            text += self._format_dimer(dimers[0], seq1, seq2)
        if self.dimerMin_3prim_dG():
            text += self._format_dimer(self._min_3prim_dimer, seq1, seq2)
        return text

    @classmethod
    def _format_hairpin(cls, hairpin, seq):
        """The folded sequence drawn as two rows joined at the loop (SecStructures.py:779-829)."""
        seq = str(seq)
        fm, rm = hairpin.fwd_matches, hairpin.rev_matches
        loop = rm[-1] - fm[-1] - 1
        odd = loop % 2
        half = (loop - odd) // 2
        fwd_tail, rev_tail = fm[0], len(seq) - rm[0] - 1
        turn = fm[-1] + half + 1                       # first base of the lower row (after the middle base, if any)
        bars = '|'
        for a, b in zip(fm, fm[1:]):
            bars += ' ' * (b - a - 1) + '|'
        shift = abs(fwd_tail - rev_tail)
        pad_top, pad_bottom = ('', ' ' * shift) if fwd_tail > rev_tail else (' ' * shift, '')
        text = "%s5' %s\n" % (pad_top, seq[:turn])
        text += '%s%s%s%s\n' % (' ' * (max(fwd_tail, rev_tail) + 3), bars, ' ' * half, seq[turn] if odd else ']')
        text += "%s3' %s\n" % (pad_bottom, seq[turn + odd:][::-1])
        if hairpin.dG is not None:
            text += 'dG(%.1fC) = %.2f kcal/mol\n' % (tdf.PCR_P.PCR_T, hairpin.dG)
        if hairpin.conversion_degree is not None:
            text += 'conversion degree = %.4f%%\n' % (hairpin.conversion_degree * 100)
        if rm[0] > len(seq) - 4:
            text += "(3'-hairpin)\n"
        return text + '\n'

    def _format_hairpins_header(self, hairpins):
        lines = ['Hairpins count: %d' % len(hairpins)]
        if self._3prim_hairpins:
            lines.append("3'-hairpins count: %d" % self._3prim_hairpins)
        lines.append('Most stable hairpin: %.2f kcal/mol' % self.hairpinMin_dG())
        if self._3prim_hairpins:
            lines.append("Most stable 3' hairpin: %.2f kcal/mol" % self.hairpinMin_3prim_dG())
        return '\n'.join(lines) + '\n\n'

    def _format_hairpins(self, hairpins, seq):
        if not hairpins:
            return 'No hairpins found\n\n'
        return self._format_hairpins_header(hairpins) + ''.join(self._format_hairpin(h, seq) for h in hairpins)

    def _format_hairpins_short(self, hairpins, seq):
        if not hairpins:
            return 'No hairpins found\n\n'
        text = self._format_hairpins_header(hairpins)
        if self.hairpinMin_dG() and hairpins[0] is not self._min_3prim_hairpin:
            text += self._format_hairpin(hairpins[0], seq)
        if self.hairpinMin_3prim_dG():
            text += self._format_hairpin(self._min_3prim_hairpin, seq)
        return text
