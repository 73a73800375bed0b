"""ORACLE (test infrastructure, see oracle/README.md) — degenerate primer.

Restates the parts of DegenPrimer/Primer.py the hot path reads: master
sequence, unambiguous variants and their ids (:243-256), per-variant
concentration (:180-186) and Tm_min/mean/max (:121-140).  Biopython's
SeqRecord/Seq are replaced by two tiny stand-ins with the attributes used.
"""
from math import log

from . import thermo
from .sequtils import reverse_complement, unambiguous_sequences


class Seq(str):
    """Just enough of Bio.Seq.Seq: str(), len(), slicing, complement()."""
    def reverse_complement(self):
        return Seq(reverse_complement(self))

    def upper(self):
        return Seq(str.upper(self))

    def __getitem__(self, k):
        r = str.__getitem__(self, k)
        return Seq(r) if isinstance(k, slice) else r


class SeqRecord(object):
    """Just enough of Bio.SeqRecord.SeqRecord."""
    def __init__(self, seq, id='', name='', description='', annotations=None):
        self.seq = seq if isinstance(seq, Seq) else Seq(seq)
        self.id = id
        self.name = name
        self.description = description
        self.annotations = annotations or {}

    def __len__(self):
        return len(self.seq)

    def upper(self):
        return SeqRecord(self.seq.upper(), self.id, self.name, self.description,
                         self.annotations)


class Primer(object):
    def __init__(self, seq_rec, concentration, generate_components=True):
        if not isinstance(seq_rec, SeqRecord):
            seq_rec = SeqRecord(seq_rec, id='primer')
        self._seq_record = seq_rec
        self._concentration = concentration
        self._unambiguous = None
        self._n_components = None
        self.Tm_max = self.Tm_mean = self.Tm_min = None
        if generate_components:
            self.generate_components()

    def __len__(self):
        return len(self._seq_record)

    def __str__(self):
        return str(self._seq_record.seq)

    def __contains__(self, seq):
        return str(seq) in self.str_sequences

    id = property(lambda self: self._seq_record.id)
    master_sequence = property(lambda self: self._seq_record)
    num_components = property(lambda self: self._n_components)
    total_concentration = property(lambda self: self._concentration)

    def generate_components(self):
        """Primer.py:107-119 + generate_unambiguous :243-256."""
        rec = self._seq_record
        variants = unambiguous_sequences(str(rec.seq))
        n = len(variants)
        out = []
        if n >= 2:
            width = int(log(n, 10) + 1)
            for i, s in enumerate(variants):
                out.append(SeqRecord(s, id='%s_%0*d' % (rec.id, width, i + 1),
                                     description=rec.description))
        self._unambiguous = out
        self._n_components = len(out) if out else 1
        return True

    @property
    def seq_records(self):
        """Primer.py:154-160."""
        return self._unambiguous if self._unambiguous else [self._seq_record]

    @property
    def str_sequences(self):
        return [str(r.seq) for r in self.seq_records]

    @property
    def concentration(self):
        """Primer.py:180-186."""
        if self._unambiguous:
            return self._concentration / float(len(self._unambiguous))
        return self._concentration

    @property
    def self_complement(self):
        s = str(self._seq_record.seq)
        return s == reverse_complement(s)

    def calculate_Tms(self):
        """Primer.py:121-140 (serial branch)."""
        if self.self_complement:
            return False
        c = self.concentration
        temps = [thermo.primer_template_Tm(s, c) for s in self.str_sequences]
        if not temps:
            return False
        self.Tm_max = max(temps)
        self.Tm_min = min(temps)
        self.Tm_mean = (self.Tm_max + self.Tm_min) / 2.0
        return True
