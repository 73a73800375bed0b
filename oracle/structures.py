"""ORACLE (test infrastructure, see oracle/README.md) — Dimer and Duplex.

Restates DegenPrimer/SecStructures.py:30-32 (thresholds), :58-192 (Dimer:
match-index structure and the terminal-loop `alternatives`), :223-387 (Duplex:
dG/K filter, stable sort, 3'-mismatch partition, ASCII rendering).
"""
from . import conditions as cond
from . import thermo
from .sequtils import reverse_complement

MAX_DIMER_DG = -3     # SecStructures.py:30
MIN_K = 100           # SecStructures.py:32


class Dimer(object):
    """SecStructures.py:58-192."""
    __slots__ = ('fwd_matches', 'offset', 'num_matches',
                 'conversion_degree', 'dG', 'K', 'fwd_mismatch')

    def __init__(self, fwd_matches=None, offset=0):
        if fwd_matches:
            self.fwd_matches = tuple(sorted(fwd_matches))
            self.offset = offset
            self.num_matches = len(self.fwd_matches)
        else:
            self.fwd_matches = ()
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

    @property
    def fwd_min(self):
        return self.fwd_matches[0]

    @property
    def fwd_max(self):
        return self.fwd_matches[-1]

    @property
    def rev_min(self):
        return self.fwd_matches[0] - self.offset

    @property
    def rev_max(self):
        return self.fwd_matches[-1] - self.offset

    @property
    def alternatives(self):
        """SecStructures.py:126-152.

        Split the matches into maximal runs of consecutive indices, then peel
        terminal runs off in the reference's order: for each left cut l, first
        every right cut (from the far right inwards), then the next left cut.
        """
        if self.num_matches < 2:
            return []
        runs = [set()]
        fm = self.fwd_matches
        for i in range(self.num_matches - 1):
            runs[-1].add(fm[i])
            if fm[i + 1] - fm[i] > 1:
                runs.append(set())
        runs[-1].add(fm[-1])
        n_loops = len(runs) - 1
        if not n_loops:
            return []
        out = []
        left = self
        for l in range(n_loops):
            right = left
            for r in range(n_loops, l, -1):
                right = Dimer(set(right.fwd_matches) - runs[r], self.offset)
                out.append(right)
            left = Dimer(set(left.fwd_matches) - runs[l], self.offset)
            out.append(left)
        return out

    @classmethod
    def from_sequences(cls, forward, revcomp, offset=0):
        """SecStructures.py:173-191 — plain character equality."""
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
    """SecStructures.py:223-387."""

    def __init__(self, fwd_seq, rev_seq, name='', dimer=None, revcomp=False):
        self.name = name
        self.fwd_len = len(fwd_seq)
        self.rev_len = len(rev_seq)
        self._fwd_sequence = fwd_seq
        self._rev_sequence = reverse_complement(rev_seq) if revcomp else rev_seq
        if not dimer:
            dimer = Dimer.from_sequences(
                fwd_seq, rev_seq if revcomp else reverse_complement(rev_seq))
        self._dimers = [dimer] + dimer.alternatives
        self._fwd_matches = []
        self._fwd_mismatches = []
        self.fwd_3_overhang = None
        self._nonzero = False
        self._recalculate()

    def __bool__(self):
        return self._nonzero

    def __hash__(self):
        return hash((self._fwd_sequence, self._rev_sequence, self._dimers[0]))

    def __eq__(self, other):
        return hash(self) == hash(other)

    def _recalculate(self):
        """SecStructures.py:294-324."""
        kept = []
        for d in self._dimers:
            if not d:
                continue
            d.dG = thermo.dimer_dG_corrected(d, self._fwd_sequence, self._rev_sequence)
            if d.dG > MAX_DIMER_DG:
                continue
            d.K = thermo.equilibrium_constant(d.dG, cond.PCR_P.PCR_T)
            if d.K < MIN_K:
                continue
            d.fwd_mismatch = d.fwd_max < self.fwd_len - 1
            kept.append(d)
        self._nonzero = bool(kept)
        if not self._nonzero:
            self._dimers = ()
            return
        self._dimers = tuple(sorted(kept, key=lambda x: x.dG))   # stable
        fwd_tail = self.fwd_len - 1 - self._dimers[0].fwd_max
        rev_tail = self.rev_len - 1 - self._dimers[0].rev_max
        self.fwd_3_overhang = fwd_tail - rev_tail if fwd_tail > rev_tail else 0
        for i, d in enumerate(self._dimers):
            if d.fwd_max < self.fwd_len - 1:
                self._fwd_mismatches.append(i)
            else:
                self._fwd_matches.append(i)

    fwd_seq = property(lambda self: self._fwd_sequence)
    rev_seq = property(lambda self: self._rev_sequence)
    dimers = property(lambda self: self._dimers)
    dimer = property(lambda self: self._dimers[0])
    dG = property(lambda self: self._dimers[0].dG)
    K = property(lambda self: self._dimers[0].K)
    have_3_matches = property(lambda self: bool(self._fwd_matches))

    @property
    def fwd_3_matches(self):
        return [self._dimers[i] for i in self._fwd_matches]

    @property
    def fwd_3_mismatches(self):
        return [self._dimers[i] for i in self._fwd_mismatches]

    @classmethod
    def print_dimer(cls, fwd_sequence, rev_sequence, dimer):
        """SecStructures.py:355-386."""
        if not dimer:
            return fwd_sequence + '\n' + rev_sequence[::-1] + '\n'
        bars = [' '] * len(fwd_sequence)
        for m in dimer.fwd_matches:
            bars[m] = '|'
        bars = ''.join(bars)
        out = ''
        if dimer.offset > 0:
            out += "5' " + fwd_sequence + " 3'\n   "
            out += bars + '\n'
            out += ' ' * dimer.offset + "3' " + rev_sequence[::-1] + " 5'\n"
        else:
            pad = ' ' * abs(dimer.offset) if dimer.offset < 0 else ''
            out += pad
            out += "5' " + fwd_sequence + " 3'\n   "
            out += pad + bars + '\n'
            out += "3' " + rev_sequence[::-1] + " 5'\n"
        if dimer.dG is not None:
            out += 'dG(%.1fC) = %.2f kcal/mol\n' % (cond.PCR_P.PCR_T, dimer.dG)
        if dimer.conversion_degree is not None:
            out += 'conversion degree = %.4f%%\n' % (dimer.conversion_degree * 100)
        return out

    def print_most_stable(self, include_fwd_3_mismatch=True):
        """SecStructures.py:279-292."""
        out = ''
        if not self._dimers:
            return out
        if not include_fwd_3_mismatch:
            if self._fwd_matches:
                return self.print_dimer(self._fwd_sequence, self._rev_sequence,
                                        self._dimers[self._fwd_matches[0]])
            out += "There's no annealing variant without 3' mismatches\n"
        return out + self.print_dimer(self._fwd_sequence, self._rev_sequence, self._dimers[0])

    def __str__(self):
        out = '%d variants of annealing sorted by stability:\n' % len(self._dimers)
        for i, d in enumerate(self._dimers):
            out += '\n%d:\n' % (i + 1)
            out += self.print_dimer(self._fwd_sequence, self._rev_sequence, d)
            out += '\n'
        return out + '\n'

    __repr__ = __str__


def structure_key(dimer):
    """The tuple hashed by SecStructures._structure_hash (DP/SecStructures.py:597-605):
    `struct.fwd_matches + struct.rev_matches` is tuple + numpy array, i.e. an ELEMENT-WISE sum."""
    return tuple(2 * f - dimer.offset for f in dimer.fwd_matches)


def find_dimers(seq1, seq2, max_dimer_dG=-3):
    """SecStructures._find_dimers, DP/SecStructures.py:609-668.

    Returns (dimers, fronts, n_3prim, min_3prim_dG): the most stable structure of every relative
    position with dG < max_dimer_dG (insertion order, de-duplicated by structure_key), the full dimer
    of every position with at least one match (first of each structure_key: the hairpin seeds), and the 3'-dimer bookkeeping of :660-667.
    """
    from . import thermo
    fwd_str = str(seq1)
    rev_str = reverse_complement(str(seq2))
    fwd_len, rev_len = len(fwd_str), len(rev_str)
    dimers, fronts, seen, seen_fronts = [], [], set(), set()
    n_3prim, min_3prim_dG = 0, None
    for front in range(fwd_len + rev_len - 1):
        fwd_left = max(fwd_len - front - 1, 0)
        fwd_right = fwd_len
        if front >= rev_len:
            fwd_right -= front - rev_len - 1
        rev_left = front - fwd_len + 1 if front >= fwd_len else 0
        rev_right = rev_len if front >= rev_len else front + 1
        matches = [f for f, r in zip(range(fwd_left, fwd_right), range(rev_left, rev_right))
                   if fwd_str[f] == rev_str[r]]
        if not matches:
            continue
        dimer = Dimer(matches, fwd_left - rev_left)
        if structure_key(dimer) not in seen_fronts:          # :643-645 (the reference does this for self-dimers only)
            seen_fronts.add(structure_key(dimer))
            fronts.append(dimer)
        min_dG, best = max_dimer_dG, None
        for d in [dimer] + dimer.alternatives:
            d.dG = thermo.dimer_dG_corrected(d, fwd_str, str(seq2))
            if min_dG > d.dG:
                min_dG, best = d.dG, d
        if best is None:
            continue
        key = structure_key(best)
        if key in seen:
            continue
        seen.add(key)
        dimers.append(best)
        if best.fwd_matches[-1] > fwd_len - 4 or best.rev_min < 3:
            n_3prim += 1
            if min_3prim_dG is None or min_3prim_dG > min_dG:
                min_3prim_dG = min_dG
    return dimers, fronts, n_3prim, min_3prim_dG


class Hairpin(object):
    """DP/SecStructures.py:194-220."""
    __slots__ = ['fwd_matches', 'rev_matches', 'conversion_degree', 'dG', 'K']

    def __init__(self, fwd_matches, rev_matches):
        if len(fwd_matches) != len(rev_matches):
            raise ValueError('number of forward matches should equal that of reverse matches.')
        self.fwd_matches = tuple(sorted(fwd_matches))
        self.rev_matches = tuple(sorted(rev_matches, reverse=True))
        self.conversion_degree = None
        self.dG = None
        self.K = None


def trim_hairpin_seed(fwd_matches):
    """The while loop of SecStructures._find_hairpins (:684-690): both sets start as the self-dimer's
    fwd_matches; pairs are dropped from the middle until the stem ends are >= 4 apart (3 unpaired bases)."""
    fwd, rev = set(fwd_matches), set(fwd_matches)
    while (rev and fwd) and (min(rev) < max(fwd) or min(rev) - max(fwd) < 4):
        fwd.remove(max(fwd))
        rev.remove(min(rev))
    return fwd, rev


def find_hairpins(seeds, seq, max_hairpin_dG=3):
    """SecStructures._find_hairpins, DP/SecStructures.py:676-716.  seeds: the full self-dimers of
    find_dimers (`fronts`).  Returns (hairpins, n_3prim, min_3prim_dG)."""
    from . import thermo
    seq = str(seq)
    seq_len = len(seq)
    hairpins, seen = [], set()
    n_3prim, min_3prim_dG = 0, None
    for seed in seeds:
        fwd, rev = trim_hairpin_seed(seed.fwd_matches)
        if not fwd:
            continue
        dimer = Dimer(fwd, min(fwd) - (seq_len - 1 - max(rev)))
        min_dG, best = max_hairpin_dG, None
        for d in [dimer] + dimer.alternatives:
            h = Hairpin(d.fwd_matches, set(seq_len - 1 - (m - d.offset) for m in d.fwd_matches))
            h.dG = thermo.hairpin_dG_corrected(h, seq)
            if min_dG > h.dG:
                min_dG, best = h.dG, h
        if best is None:
            continue
        key = best.fwd_matches + best.rev_matches          # _structure_hash: tuple + tuple = concatenation
        if key in seen:
            continue
        seen.add(key)
        hairpins.append(best)
        if best.rev_matches[0] > seq_len - 4:
            n_3prim += 1
            if min_3prim_dG is None or min_3prim_dG > min_dG:
                min_3prim_dG = min_dG
    return hairpins, n_3prim, min_3prim_dG
