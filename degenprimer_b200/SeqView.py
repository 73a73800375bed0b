"""Template loader: FASTA / GenBank / EMBL files, or a packed cache -> records resident in HBM as 2-bit planes.

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


def parse_genbank(path):
    """Records of a GenBank flat file as (id, description, sequence): id = VERSION (accession.version) if
    present, else the first ACCESSION, else the LOCUS name — the order Bio.SeqIO's 'gb' parser uses, which
    the reference reads these files with (BioUtils/SeqUtils.py:30-35,160-172); description = DEFINITION
    joined to one line; sequence = the ORIGIN block without digits and blanks."""
    out, cur, field = [], None, None
    with open(path) as fh:
        for line in fh:
            key = line[:12].strip()
            if line.startswith('LOCUS'):
                parts = line.split()
                cur = {'locus': parts[1] if len(parts) > 1 else '', 'acc': '', 'ver': '', 'def': [], 'seq': []}
                field = None
            elif cur is None:
                continue
            elif line.startswith('//'):
                rid = cur['ver'] or cur['acc'] or cur['locus']
                out.append((rid, ' '.join(cur['def']), ''.join(cur['seq'])))
                cur, field = None, None
            elif field == 'ORIGIN':
                cur['seq'].append(''.join(c for c in line if c.isalpha() or c == '*'))
            elif key == 'DEFINITION':
                cur['def'].append(line[12:].strip())
                field = 'DEFINITION'
            elif key == 'ACCESSION':
                cur['acc'] = (line[12:].split() or [''])[0]
                field = None
            elif key == 'VERSION':
                cur['ver'] = (line[12:].split() or [''])[0]
                field = None
            elif key == 'ORIGIN':
                field = 'ORIGIN'
            elif key == '' and field == 'DEFINITION' and line[:12].strip() == '':
                cur['def'].append(line.strip())
            elif key:
                field = None
    if cur is not None:
        raise ValueError('%s: GenBank record %s is not terminated by //' % (path, cur['locus']))
    return out


def parse_embl(path):
    """Records of an EMBL flat file as (id, description, sequence): id = first AC accession plus the
    sequence version of the ID line (SV n) when there is one, else the ID name; description = DE lines
    joined; sequence = the lines after SQ without digits and blanks."""
    out, cur, in_seq = [], None, False
    with open(path) as fh:
        for line in fh:
            tag = line[:2]
            if tag == 'ID':
                fields = [f.strip() for f in line[5:].split(';')]
                sv = next((f[3:].strip() for f in fields if f.startswith('SV ')), '')
                cur = {'name': fields[0].split()[0] if fields and fields[0] else '', 'sv': sv, 'acc': '', 'de': [],
                       'seq': []}
                in_seq = False
            elif cur is None:
                continue
            elif tag == '//':
                rid = cur['acc'] or cur['name']
                if cur['acc'] and cur['sv']:
                    rid = '%s.%s' % (cur['acc'], cur['sv'])
                out.append((rid, ' '.join(cur['de']), ''.join(cur['seq'])))
                cur, in_seq = None, False
            elif in_seq:
                cur['seq'].append(''.join(c for c in line if c.isalpha() or c == '*'))
            elif tag == 'AC' and not cur['acc']:
                cur['acc'] = line[5:].split(';')[0].strip()
            elif tag == 'DE':
                cur['de'].append(line[5:].strip())
            elif tag == 'SQ':
                in_seq = True
    if cur is not None:
        raise ValueError('%s: EMBL record %s is not terminated by //' % (path, cur['name']))
    return out


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


def source_identity(path):
    """What a packed cache remembers about the file it was made from."""
    st = os.stat(path)
    return '%s|%d|%d' % (os.path.abspath(path), st.st_size, st.st_mtime_ns)


def _cache_origin(cache):
    try:
        return nat.packed_info(cache)['origin']
    except (ValueError, OSError):
        return None


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
        schema = next(iter(schemas))
        ctx = nat.context(device)
        for i, f in enumerate(valid):
            # `dbname` is the reference's on-disk index (SeqIO.index_db, BioUtils/SeqUtils.py:160-172);
            # here it names the packed-template cache.  The cache records what it was
            # packed from (absolute path, size, mtime in ns -- inside the file, under its checksum) and
            # is used only when that still describes the source file; SeqIO.index_db likewise refuses an
            # index whose file list differs (BioUtils/SeqUtils.py:169).  Anything else is rebuilt.
            cache = None if not dbname else (dbname if len(valid) == 1 else '%s.%d' % (dbname, i))
            origin = source_identity(f)
            if cache and os.path.isfile(cache) and _cache_origin(cache) == origin:
                t = ctx.template_load(cache)
            else:
                if schema == 'fasta':
                    t = ctx.template_from_fasta(f)
                else:                       # GenBank / EMBL: parsed on the host, packed on the device
                    recs = sorted(parse_genbank(f) if schema == 'gb' else parse_embl(f))
                    if not recs:
                        raise ValueError('no %s records in %s' % (schema, f))
                    for (a, _, _), (b, _, _) in zip(recs, recs[1:]):
                        if a == b:
                            raise ValueError("Duplicate key '%s'" % a)
                    t = ctx.template_from_sequences([r[2] for r in recs])
                    t.set_record_names([r[0] for r in recs], [r[1] for r in recs])
                if cache:
                    t.origin = origin
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
