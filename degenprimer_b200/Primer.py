"""Degenerate primer — host mirror of DegenPrimer/Primer.py:36-289.

Same constructor, properties and variant ids; the per-variant melting
temperatures (`calculate_Tms`, Primer.py:121-140) are computed by one launch of
the Tm kernel instead of a Python loop / process pool.
"""
from math import log

from . import TD_Functions as tdf
from .SeqUtils import IUPAC_ambiguous, Seq, SeqRecord, unambiguous_sequences


class Primer(object):
    def __init__(self, seq_rec, concentration, generate_components=False):
        if isinstance(seq_rec, str):
            seq_rec = SeqRecord(Seq(seq_rec.upper()), id='primer')
        self._seq_record = seq_rec
        self._concentration = concentration
        self._n_components = None
        self._unambiguous = None
        self._initialized = False
        self._Tm_max = self._Tm_mean = self._Tm_min = None
        if generate_components:
            self.generate_components()
            self.calculate_Tms()

    def _check_initialized(self):
        assert self._initialized, '%s primer was not initialized' % self._seq_record.id

    def _check_components(self):
        assert self._n_components, '%s primer was not initialized' % self._seq_record.id

    def __len__(self):
        return len(self._seq_record)

    def __contains__(self, seq):
        return str(seq) in self.str_sequences

    def __bool__(self):
        return len(self._seq_record.seq) != 0

    def __str__(self):
        return str(self._seq_record.seq)

    def __hash__(self):
        return hash(str(self._seq_record.seq))

    def __repr__(self):
        self._check_initialized()
        rep = "%s:    %db    5'-%s-3'\n" % (self.id, len(self), str(self))
        if self._unambiguous:
            rep += '   Tm max  = %.1f\n   Tm mean = %.1f\n   Tm min  = %.1f\n' % \
                   (self._Tm_max, self._Tm_mean, self._Tm_min)
            rep += '   Unambiguous primers:   %d \n' % self._n_components
        else:
            rep += '   Tm:       %.1f\n' % self._Tm_max
        return rep

    @property
    def id(self):
        return self._seq_record.id

    @property
    def num_components(self):
        return self._n_components

    def generate_components(self):
        self._unambiguous = self.generate_unambiguous(self._seq_record)
        self._n_components = len(self._unambiguous) if self._unambiguous else 1
        return True

    def calculate_Tms(self, abort_event=None):
        self._check_components()
        if self.self_complement:
            return False
        temperatures = tdf.primer_template_Tms(self.sequences, self.concentration)
        if not len(temperatures):
            return False
        self._Tm_max = float(max(temperatures))
        self._Tm_min = float(min(temperatures))
        self._Tm_mean = (self._Tm_max + self._Tm_min) / 2.0
        self._initialized = True
        return True

    @property
    def master_sequence(self):
        return self._seq_record

    @property
    def all_seq_records(self):
        self._check_components()
        return [self._seq_record] + (self._unambiguous or [])

    @property
    def seq_records(self):
        self._check_components()
        return self._unambiguous if self._unambiguous else [self._seq_record]

    @property
    def sequences(self):
        return [r.seq for r in self.seq_records]

    @property
    def str_sequences(self):
        return [str(s) for s in self.sequences]

    @property
    def total_concentration(self):
        return self._concentration

    @total_concentration.setter
    def total_concentration(self, C):
        self._concentration = C
        self.calculate_Tms()

    @property
    def concentration(self):
        self._check_components()
        if self._unambiguous:
            return self._concentration / float(len(self._unambiguous))
        return self._concentration

    Tm_max = property(lambda self: self._Tm_max)
    Tm_mean = property(lambda self: self._Tm_mean)
    Tm_min = property(lambda self: self._Tm_min)

    @property
    def self_complement(self):
        s = str(self._seq_record.seq)
        return s == str(Seq(s).reverse_complement())

    def has_subsequence(self, seq):
        return any(s.count(str(seq)) > 0 for s in self.str_sequences)

    def find_sequences(self, seq):
        return [s for s in self.sequences if str(s).count(str(seq)) > 0]

    def find_records(self, seq):
        return [r for r in self.seq_records if str(r.seq).count(str(seq)) > 0]

    def find_id(self, seq):
        for r in self.seq_records:
            if str(r.seq) == str(seq):
                return r.id
        return ''

    @classmethod
    def generate_unambiguous(cls, seq_rec):
        """Variant records with ids '<id>_<NN>' (Primer.py:243-256)."""
        variants = unambiguous_sequences(seq_rec.seq)
        n = len(variants)
        if n < 2:
            return []
        width = int(log(n, 10) + 1)
        return [SeqRecord(Seq(s), description=seq_rec.description,
                          id='%s_%0*d' % (seq_rec.id, width, i + 1))
                for i, s in enumerate(variants)]

    @classmethod
    def from_sequences(cls, seq_records, concentration, _id=None, description=None):
        """Primer.py:259-289: compose the master from unambiguous records."""
        if not seq_records:
            return None
        n = len(seq_records[0].seq)
        if any(len(r.seq) != n for r in seq_records):
            return None
        sets = [set() for _ in range(n)]
        for r in seq_records:
            for i in range(n):
                sets[i].update(unambiguous_sequences(r.seq[i].upper()))
        letters = []
        for s in sets:
            if len(s) == 1:
                letters.append(next(iter(s)))
                continue
            for letter, exp in IUPAC_ambiguous.items():
                if s == set(exp):
                    letters.append(letter)
                    break
        master = SeqRecord(Seq(''.join(letters)),
                           description=description if description else seq_records[0].description,
                           id=_id if _id else seq_records[0].id)
        return cls(master, concentration)
