"""Host-side IUPAC helpers with the reference's names.

Mirrors DegenPrimer/SeqUtils.py:30-60 (`IUPAC_ambiguous`, `unambiguous_sequences`
— the expansion ORDER defines variant ids) and the complement map of
DegenPrimer/SecStructures.py:34-55.  `Seq` / `SeqRecord` are minimal stand-ins
for the Biopython classes the reference passes around (Biopython is not a
dependency of this package): only the attributes the hot path and its callers
read are provided (SURVEY §8b "object contracts").
"""

import re

IUPAC_ambiguous = {'R': ['A', 'G'], 'Y': ['C', 'T'], 'S': ['G', 'C'], 'W': ['A', 'T'],
                   'K': ['G', 'T'], 'M': ['A', 'C'], 'B': ['C', 'G', 'T'], 'D': ['A', 'G', 'T'],
                   'H': ['A', 'C', 'T'], 'V': ['A', 'C', 'G'], 'N': ['A', 'T', 'C', 'G']}
IUPAC_unambiguous = ['A', 'T', 'G', 'C']

rc_map = {'A': 'T', 'T': 'A', 'G': 'C', 'C': 'G', 'R': 'Y', 'Y': 'R', 'S': 'S', 'W': 'W',
          'K': 'M', 'M': 'K', 'B': 'V', 'V': 'B', 'D': 'H', 'H': 'D', 'N': 'N'}
_RC_TABLE = str.maketrans(''.join(rc_map.keys()), ''.join(rc_map.values()))
_ALPHABET = frozenset(rc_map)


def reverse_complement(seq):
    """DegenPrimer/SecStructures.py:50-55: upper-cases; KeyError on foreign letters."""
    s = str(seq).upper()
    bad = set(s) - _ALPHABET
    if bad:
        raise KeyError(sorted(bad)[0])
    return s.translate(_RC_TABLE)[::-1]


def unambiguous_sequences(s):
    """All unambiguous expansions; the earliest degenerate position varies fastest
    (DegenPrimer/SeqUtils.py:44-60)."""
    out = ['']
    for letter in str(s):
        repl = IUPAC_ambiguous.get(letter)
        n = len(out)
        if repl is None:
            out = [x + letter for x in out]
        else:
            out = [out[i % n] + repl[i // n] for i in range(n * len(repl))]
    return out


class Seq(str):
    """str with the few Bio.Seq.Seq methods the reference calls."""
    __slots__ = ()

    def reverse_complement(self):
        return Seq(reverse_complement(self))

    def complement(self):
        return Seq(reverse_complement(self)[::-1])

    def upper(self):
        return Seq(str.upper(self))

    def __getitem__(self, k):
        r = str.__getitem__(self, k)
        return Seq(r) if isinstance(k, slice) else r


class SeqRecord(object):
    """seq / id / name / description / annotations, len(), upper(), slicing."""

    def __init__(self, seq, id='<unknown id>', name='<unknown name>',
                 description='<unknown description>', annotations=None):
        self.seq = seq if isinstance(seq, Seq) or not isinstance(seq, str) else Seq(seq)
        self.id = id
        self.name = name
        self.description = description
        self.annotations = annotations if annotations is not None else {}

    def __len__(self):
        return len(self.seq)

    def __getitem__(self, k):
        if isinstance(k, slice):
            return SeqRecord(self.seq[k], self.id, self.name, self.description, dict(self.annotations))
        return self.seq[k]

    def upper(self):
        return SeqRecord(self.seq.upper(), self.id, self.name, self.description, dict(self.annotations))

    def __repr__(self):
        return 'SeqRecord(id=%r, len=%d)' % (self.id, len(self))


def load_sequence(seq_string, rec_id='', desc=''):
    """Raw-string branch of DegenPrimer/SeqUtils.py:62-112 (file input: use SeqView)."""
    sequence = seq_string.replace(' ', '').upper()
    if not sequence or set(sequence) - _ALPHABET:
        raise ValueError('Primer.load_sequence: given string is not a valid sequence: %s' % seq_string)
    return SeqRecord(Seq(sequence), id=rec_id, name=desc, description=desc)


_ILLEGAL = re.compile(r'[=/]')


def pretty_rec_name(rec):
    """BioUtils/BioUtils/SeqUtils.py:536-543: annotation 'source', else the
    description without the id, else the id; '=' and '/' become '-'."""
    if rec.description is None:
        rec.description = ''
    name = (rec.annotations.get('source')
            or rec.description.replace(rec.id, '').strip()
            or rec.id)
    return _ILLEGAL.sub('-', name)
