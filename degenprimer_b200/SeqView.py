"""Template loader: FASTA files -> records resident in HBM as 2-bit planes.

Keeps the container interface of BioUtils/BioUtils/SeqUtils.py:103-271
(`SeqView`: `load`, `load_dir`, `keys`, `__getitem__` by index / id / slice,
`subview`, `get`, `close`, iteration) that DegenPrimer/iPCR_Base.py:47-52 and
DegenPrimer/iPCR.py:68-118 use, but a record is a handle into device memory
instead of a Biopython SeqRecord parsed from an SQLite index: the sequence is
packed once by dp_template_from_fasta and never copied on the host again.
"""
import os
import re

from . import _native as nat
from .SeqUtils import Seq

_FASTA = re.compile(r'.*\.f(asta|as|a|aa|fn|rn|na)$')     # SeqLoader schemas, SeqUtils.py:30-35
_OTHER = re.compile(r'.*\.(gb|gbk|embl)$')


class StrandView(object):
    """Read-only, str-like view of one strand of a resident record.

    Stands in for the `fwd_seq` / `rev_seq` strings of the 7-tuple that
    SearchEngine.find_matches returns (DegenPrimer/SearchEngine.py:325,359):
    len(), slicing and str() work; text is decoded from HBM on demand.
    """
    __slots__ = ['template', 'record', 'reverse', '_text']

    def __init__(self, template, record, reverse=False, text=None):
        self.template, self.record, self.reverse, self._text = template, record, reverse, text

    def __len__(self):
        return self.template.lengths[self.record]

    def __str__(self):
        if self._text is None:
            self._text = self.template.unpack(self.record, 0, None, self.reverse)
        return self._text

    def __getitem__(self, k):
        if self._text is not None:
            return self._text[k]
        n = len(self)
        if isinstance(k, slice):
            start, stop, step = k.indices(n)
            if step == 1:
                return self.template.unpack(self.record, start, max(0, stop - start), self.reverse)
            return str(self)[k]
        if k < 0:
            k += n
        if not 0 <= k < n:
            raise IndexError('string index out of range')
        return self.template.unpack(self.record, k, 1, self.reverse)

    def __eq__(self, other):
        return str(self) == str(other)

    def __hash__(self):
        return hash(str(self))

    def upper(self):
        return self

    def reverse_complement(self):
        return StrandView(self.template, self.record, not self.reverse)

    def __reduce__(self):
        return (Seq, (str(self),))


class ResidentRecord(object):
    """SeqRecord duck type whose sequence lives on the device."""

    def __init__(self, template, record, annotations=None):
        self._template = template
        self._record = record
        self.id = template.ids[record]
        self.name = self.id
        self.description = template.descriptions[record]
        self.annotations = annotations if annotations is not None else {}
        self.seq = StrandView(template, record)

    def __len__(self):
        return self._template.lengths[self._record]

    def upper(self):
        return self

    def __repr__(self):
        return 'ResidentRecord(id=%r, len=%d)' % (self.id, len(self))


class SeqView(object):
    def __init__(self, upper=False):
        self.upper = upper           # kept for signature parity: input is always case-folded
        self._templates = []
        self._where = {}             # id -> (template, record index)
        self._ids = ()
        self.master = False

    @staticmethod
    def guess_schema(filename):
        f = filename.lower()
        if _FASTA.match(f):
            return 'fasta'
        m = _OTHER.match(f)
        return ('gb' if m.group(1) in ('gb', 'gbk') else 'embl') if m else None

    def load_dir(self, dirname, dbname=None, namefilter=None):
        if isinstance(namefilter, str):
            namefilter = re.compile(namefilter).match
        elif hasattr(namefilter, 'match'):
            namefilter = namefilter.match
        elif not callable(namefilter):
            namefilter = lambda n: True
        files = [os.path.join(dirname, f) for f in sorted(os.listdir(dirname)) if namefilter(f)]
        files = [f for f in files if os.path.isfile(f)]
        if not files:
            print('No files found in: %s' % dirname)
            return False
        return self.load(files, dbname)

    def load(self, files, dbname=None, device=None):
        if isinstance(files, (str, bytes)):
            files = [files]
        self.close()
        valid, schemas = [], set()
        for f in files:
            if isinstance(f, bytes):
                f = f.decode('utf8')
            if not os.path.isfile(f):
                print('No such file: %s' % f)
                continue
            schema = self.guess_schema(f)
            if not schema:
                print('Unable to guess schema from filename: %s' % f)
                continue
            schemas.add(schema)
            valid.append(f)
        if len(schemas) != 1:
            raise ValueError('All files should be of the same type, but %d types found: %s'
                             % (len(schemas), ', '.join(sorted(schemas))))
        if schemas != {'fasta'}:
            raise NotImplementedError('only FASTA templates are packed on the device; GenBank/EMBL '
                                      'input is listed under "next" (SURVEY §8f N4)')
        ctx = nat.context(device)
        for i, f in enumerate(valid):
            # `dbname` is the reference's on-disk index (SeqIO.index_db, BioUtils/SeqUtils.py:160-172);
            # here it names the packed-template cache: used when it is newer than the
            # FASTA file, (re)written otherwise.
            cache = None if not dbname else (dbname if len(valid) == 1 else '%s.%d' % (dbname, i))
            if cache and os.path.isfile(cache) and os.path.getmtime(cache) >= os.path.getmtime(f):
                t = ctx.template_load(cache)
            else:
                t = ctx.template_from_fasta(f)
                if cache:
                    t.save(cache)
            self._templates.append(t)
            for r, rid in enumerate(t.ids):
                if rid in self._where:
                    raise ValueError("Duplicate key '%s'" % rid)
                self._where[rid] = (t, r)
        self._ids = tuple(sorted(self._where))
        self.master = True
        return bool(self)

    def load_packed(self, paths, device=None):
        """Packed-template files written by `save_packed` / `load(..., dbname=...)`: the planes go
        straight to HBM, nothing is parsed (dp_template_load)."""
        if isinstance(paths, (str, bytes)):
            paths = [paths]
        self.close()
        ctx = nat.context(device)
        for f in paths:
            t = ctx.template_load(f)
            self._templates.append(t)
            for r, rid in enumerate(t.ids):
                if rid in self._where:
                    raise ValueError("Duplicate key '%s'" % rid)
                self._where[rid] = (t, r)
        self._ids = tuple(sorted(self._where))
        self.master = True
        return bool(self)

    def save_packed(self, path):
        """Write the resident templates as packed files; returns the list of paths
        (`path` for one template, `path.0`, `path.1`, ... for several)."""
        one = len(self._templates) == 1
        return [t.save(path if one else '%s.%d' % (path, i)) for i, t in enumerate(self._templates)]

    @classmethod
    def safe_load(cls, files):
        try:
            view = cls()
            view.load(files)
        except Exception as e:       # same contract as SeqUtils.py:176-189
            print(str(e))
            return None
        if len(view) == 0:
            print('No sequences were loaded from:\n%s' % files)
            return None
        return view

    def close(self):
        if self.master:
            for t in self._templates:
                t.close()
        self._templates, self._where, self._ids, self.master = [], {}, (), False

    def __bool__(self):
        return bool(self._ids)

    def __len__(self):
        return len(self._ids)

    def __iter__(self):
        return (self[k] for k in self._ids)

    def __getitem__(self, key):
        if isinstance(key, int):
            return self[self._ids[key]]
        if isinstance(key, str):
            if key not in self._where:
                raise KeyError('Sequence %s not found' % key)
            t, r = self._where[key]
            return ResidentRecord(t, r)
        if isinstance(key, slice):
            return self.subview(self._ids[key])
        raise KeyError('Sequence %s not found' % (key,))

    def get(self, key, default=None):
        try:
            return self[key]
        except Exception:
            return default

    def keys(self):
        return self._ids

    iterkeys = lambda self: iter(self._ids)

    def key(self, index):
        return self._ids[index]

    def values(self, keys=None):
        return [self[k] for k in (keys or self._ids)]

    def items(self):
        return {k: self[k] for k in self._ids}

    def subview(self, keys):
        v = SeqView(self.upper)
        v._templates = self._templates
        v._where = self._where
        v._ids = tuple(keys)
        v.master = False
        return v
