"""Region / Product — host mirror of DegenPrimer/Product.py:30-215.

Plain host objects: a PCR product is a template region bounded by the 3' ends of the forward and
reverse primers that produce it, plus the sets of annealed primers (Duplex objects) on either side.
PCR_Simulation and the reporters of the reference read `start, end, name, quantity, cycles,
fwd_primers, rev_primers, fwd_template, rev_template, fwd_ids, rev_ids, pretty_print()`.
"""
from copy import deepcopy
from math import log

TEXT_WIDTH = 80      # BioUtils/Tools/Text.py:16


def wrap_text(text, width=TEXT_WIDTH):
    """Greedy word wrap of BioUtils/Tools/Text.py:33-55: runs of blanks collapse, a line holds at most
    width-1 characters, a word longer than `width` is cut into hyphenated pieces; ends with a newline."""
    limit = width - 1
    lines, cur = [], ''
    for word in text.split():
        if cur and len(cur) + len(word) > limit:
            lines.append(cur)
            cur = ''
        while not cur and len(word) > width:
            lines.append(word[:limit] + '-')
            word = word[limit:]
        cur = cur + ' ' + word if cur else word
    lines.append(cur)
    return '\n'.join(lines) + '\n'


def hr(text='', symbol='-', width=TEXT_WIDTH):
    """Centered horizontal rule (BioUtils/Tools/Text.py:22-26): an empty caption gives just the blank
    lines; captions longer than the line are wrapped to 3/4 of it first."""
    if len(text) > TEXT_WIDTH - 4:
        lines = wrap_text(text, TEXT_WIDTH * 3 // 4).split('\n')
    else:
        lines = [text]
    return '\n'.join((' %s ' % l).center(width, symbol) for l in lines if l) + '\n\n'


def format_concentration(quantity, unit='M'):
    """TD_Functions.format_concentration = Text.format_quantity(c, 'M') (BioUtils/Tools/Text.py:125-139)."""
    if quantity is None or quantity < 0:
        return 'N/A  %s' % unit
    if quantity == 0:
        return '0.0  %s' % unit
    mag = -log(quantity, 10)
    for top, scale, prefix in ((0, 1, ' '), (3, 1e3, 'm'), (6, 1e6, 'u'), (9, 1e9, 'n'), (12, 1e12, 'p'),
                               (15, 1e15, 'f')):
        if mag < top:
            return '%.1f %s%s' % (quantity * scale, prefix, unit)
    return '%.1f a%s' % (quantity * 1e18, unit)


class Region(object):
    """A stretch [start, end] (1-based, inclusive) of one strand of a named sequence (Product.py:30-86)."""
    __slots__ = ['name', 'start', 'end', 'forward']

    def __init__(self, name, start, end, forward):
        if start < 1 or end < 1:
            raise ValueError('Region: start and end of a sequence region should be grater than 1.')
        if start > end:
            raise ValueError("Region: start of a sequence region should be less than it's end")
        self.name, self.start, self.end, self.forward = name, start, end, forward

    def pretty_print(self, with_name=True):
        rows = [('target', self.name)] if with_name else []
        rows += [('strand', 'forward' if self.forward else 'reverse'), ('start', '%d' % self.start),
                 ('end', '%d' % self.end), ('length', '%d' % len(self))]
        return ''.join('%-8s%s\n' % (k + ':', v) for k, v in rows)

    def __str__(self):
        return self.pretty_print()

    def __len__(self):
        return self.end - self.start + 1

    def __hash__(self):
        return hash((self.name, self.start, self.end, self.forward))

    def __eq__(self, other):
        return hash(self) == hash(other)

    def __ne__(self, other):
        return hash(self) != hash(other)

    def __iadd__(self, other):
        """Grow to the union with a region of the same sequence and strand (:71-77)."""
        if self.name == other.name and self.forward == other.forward:
            self.start = min(self.start, other.start)
            self.end = max(self.end, other.end)
        return self

    def overlaps(self, other):
        """:79-86.  The reference's expression is `same_name and same_strand and start_inside or end_inside`:
        the second alternative does not look at name or strand.  Kept as it is."""
        return ((self.name == other.name and self.forward == other.forward
                 and self.start <= other.start <= self.end)
                or self.start <= other.end <= self.end)

    def __getstate__(self):
        return tuple(getattr(self, s) for s in Region.__slots__)

    def __setstate__(self, state):
        for s, v in zip(Region.__slots__, state):
            setattr(self, s, v)


class Product(Region):
    """One PCR product (Product.py:89-215)."""
    __slots__ = ['quantity', 'cycles', 'fwd_primers', 'rev_primers', '_fwd_margin', '_rev_margin',
                 'fwd_template', 'rev_template']

    def __init__(self, template_name, start, end, fwd_primers=(), rev_primers=()):
        Region.__init__(self, template_name, start, end, forward=True)
        self.quantity = 0
        self.cycles = 0
        self.fwd_primers = set()
        self.rev_primers = set()
        self._fwd_margin = max(self.start - 1, 1)
        self._rev_margin = self.end + 1
        self.fwd_template = Region(template_name, self._fwd_margin, self._fwd_margin, forward=True)
        self.rev_template = Region(template_name, self._rev_margin, self._rev_margin, forward=False)
        for p in fwd_primers:
            self.add_fwd_primer(p)
        for p in rev_primers:
            self.add_rev_primer(p)

    def add_fwd_primer(self, primer):
        """The annealing site of a forward primer ends right before the product (:201-207)."""
        if primer in self.fwd_primers:
            return
        self.fwd_primers.add(primer)
        self.fwd_template += Region(self.name, max(self.start - primer.fwd_len, 1), max(self.start - 1, 1),
                                    forward=True)

    def add_rev_primer(self, primer):
        if primer in self.rev_primers:
            return
        self.rev_primers.add(primer)
        self.rev_template += Region(self.name, self.end + 1, self.end + primer.fwd_len, forward=False)

    def __iadd__(self, other):
        """Merge a product found again (:165-186): on either side the wider end replaces margin, template
        and primers (deep copies), an equal end pools the primers."""
        if (self.name, self.forward) != (other.name, other.forward):
            return self
        if other.start < self.start:
            self.start, self._fwd_margin = other.start, other._fwd_margin
            self.fwd_template, self.fwd_primers = deepcopy(other.fwd_template), deepcopy(other.fwd_primers)
        elif other.start == self.start:
            for p in other.fwd_primers:
                self.add_fwd_primer(p)
        if other.end > self.end:
            self.end, self._rev_margin = other.end, other._rev_margin
            self.rev_template, self.rev_primers = deepcopy(other.rev_template), deepcopy(other.rev_primers)
        elif other.end == self.end:
            for p in other.rev_primers:
                self.add_rev_primer(p)
        return self

    # Region defines __eq__, so the hash has to be named again for the subclass
    __hash__ = Region.__hash__

    @property
    def fwd_ids(self):
        return sorted(p.name for p in self.fwd_primers)

    @property
    def rev_ids(self):
        return sorted(p.name for p in self.rev_primers)

    @staticmethod
    def _rep_primers(primers, tostr=repr):
        return ''.join('%s:\n%s\n' % (p.name, tostr(p)) for p in sorted(primers, key=lambda x: x.name))

    def _status_lines(self):
        return ('concentration:    %s\n' % format_concentration(self.quantity)
                + 'number of cycles: %d\n' % self.cycles)

    def pretty_print(self, with_name=True, include_fwd_3_mismatch=True):
        """Report layout of Product.py:128-147: region, status, both annealing sites, both primer lists."""
        def show(duplex):
            return duplex.print_most_stable(include_fwd_3_mismatch)
        parts = [Region.pretty_print(self, with_name=with_name), self._status_lines()]
        for caption, site in ((' forward annealing site ', self.fwd_template),
                              (' reverse annealing site ', self.rev_template)):
            parts.append(hr(caption) + site.pretty_print(with_name=False))
        for caption, primers in ((' forward primers ', self.fwd_primers), (' reverse primers ', self.rev_primers)):
            parts.append(hr(caption) + self._rep_primers(primers, show))
        return '\n'.join(parts)

    def __str__(self):
        """Plain layout of Product.py:149-162."""
        return ''.join((Region.pretty_print(self), self._status_lines(),
                        '\nforward annealing site:\n', str(self.fwd_template),
                        '\nreverse annealing site:\n', str(self.rev_template),
                        '\nforward primers:\n', self._rep_primers(self.fwd_primers),
                        '\nreverse primers:\n', self._rep_primers(self.rev_primers)))

    def __getstate__(self):
        return Region.__getstate__(self) + tuple(getattr(self, s) for s in Product.__slots__)

    def __setstate__(self, state):
        Region.__setstate__(self, state[:4])
        for s, v in zip(Product.__slots__, state[4:]):
            setattr(self, s, v)
