"""Run the UNMODIFIED DegenPrimer sources under Python 3 — TEST INFRASTRUCTURE ONLY.

The reference (allista/DegenPrimer @ c610551c, read-only at /root/reference) is
Python-2.7-only and needs Biopython 1.66; this image has neither `python2`,
`lib2to3` nor `Bio`.  This module imports the reference's hot-path modules
anyway: a meta-path finder reads every `DegenPrimer.*` / `BioUtils.*` source
file WHERE IT LIES under /root/reference (nothing is copied into this repo),
rewrites Python-2 syntax in memory, and restores Python-2 semantics with an AST
pass before compiling.  `tests/golden/make_reference_golden.py` (scan, duplexes, dG, K,
Tm), `make_reference_dimers.py` (SecStructures: self-/cross-dimers, hairpins) and
`make_reference_mixture.py` (MixtureFactory: products) use it to record the reference's
own outputs as committed fixtures; `tests/test_reference_*.py` pin the oracle, the host
mirrors and the CUDA path to them.  Nothing under degenprimer_b200/ may import this file, and it
cannot run on the GPU box (/root/reference does not exist there).

What is rewritten (the complete list; everything else is the reference's code):

  text  (T1) `print ...` statements -> print(...)            (T2) `except E, e:` -> `except E as e:`
        (T3) `lambda(x):` -> `lambda x:`                     (T4) `exec S in D` -> exec(S, D)
        (T5) stdlib renames: Queue->queue, StringIO->io, cPickle->pickle, collections.{Sequence,Sized,Set}
             -> collections.abc, multiprocessing.queues.Queue -> a subclass that passes the fork context (3.x wants ctx=)
        (T6) `.iteritems()/.itervalues()/.iterkeys()` -> items/values/keys; `x.next()` -> `x.__next__()`
        (T7) `from array import array` -> a wrapper that takes str typecodes' initialisers as latin-1 bytes
             (py2 `array('b', 'ATGC')`, DP/SearchEngine.py:42,203)
        (T8) `open(name, 'rb')` feeding csv.reader -> text mode (DP/UnifiedNN.py:174)
        (T9) implicit relative imports of a sibling module (`from Output import ...`) -> explicit `from .Output`
  AST   (A1) every `a / b` and `a /= b` -> py2 division: floor for two integers, true division otherwise
             (DP/SearchEngine.py:216-226,254,300,312,341 rely on integer division)
        (A2) a class defining `__nonzero__` also gets `__bool__` (DP/SecStructures.py:103,257: `if dup:`)
        (A3) class-body `__metaclass__ = M` -> `class C(..., metaclass=M)` (DP/PCR_Parameters.py:67)
        (A4) `d.keys()/values()/items()` -> list(...)  (DP/UnifiedNN.py:214 mutates while iterating)
        (A5) a class defining `__div__` / `__idiv__` also answers `/` and `/=` (DP/UnifiedNN.py:65-74: Pars)
  names (N1) module globals `map`, `filter`, `zip`, `range` return lists; `xrange`, `basestring`, `unicode`,
             `long`, `reduce`, `raw_input` exist.
  deps  (D1) Biopython stand-ins `Bio.Seq.Seq`, `Bio.SeqRecord.SeqRecord`, `Bio.Alphabet.IUPAC`, `Bio.SeqIO`
             (FASTA only): immutable string wrappers with the handful of methods the path calls
             (`str`, `len`, slicing, `upper`, `complement`, `reverse_complement`, `.id/.name/.description`).

Known residual differences from a real CPython-2.7 run, none of which reaches the recorded values:
dict/set iteration order (the path's results do not depend on it: component order comes from list
arithmetic in DP/SeqUtils.py:44-60), `round()`/`repr(float)` (not used on the path), scipy 1.x instead of
0.16 `fftpack` (noise 1e-14 against a 1/6 decision margin).
"""
import ast
import builtins
import functools
import importlib.abc
import importlib.util
import os
import re
import sys
import types

REFERENCE_ROOT = os.environ.get('DEGENPRIMER_REFERENCE', '/root/reference')
_ROOTS = {
    'DegenPrimer': os.path.join(REFERENCE_ROOT, 'DegenPrimer'),
    'BioUtils': os.path.join(REFERENCE_ROOT, 'BioUtils', 'BioUtils'),
}


def available():
    return all(os.path.isdir(p) for p in _ROOTS.values())


# --------------------------------------------------------------------------- text pass
_PRINT_RE = re.compile(r'^(?P<head>\s*(?:(?:if|elif|else|for|while|try|except|finally|with)\b[^\n]*?:\s*)?)'
                       r'print(?P<chev>\s*>>\s*[^,\n]+,)?(?=[\s\'"]|$)(?P<rest>[^\n]*)$')


def _logical_lines(src):
    """Join physical lines that end in a backslash or sit inside open brackets / triple quotes."""
    out, buf, depth, triple = [], [], 0, None
    for line in src.split('\n'):
        buf.append(line)
        i, n = 0, len(line)
        in_str = None
        while i < n:
            c = line[i]
            if triple:
                if line.startswith(triple, i):
                    triple = None
                    i += 3
                    continue
            elif in_str:
                if c == '\\':
                    i += 1
                elif c == in_str:
                    in_str = None
            elif c == '#':
                break
            elif line.startswith("'''", i) or line.startswith('"""', i):
                triple = line[i:i + 3]
                i += 3
                continue
            elif c in '\'"':
                in_str = c
            elif c in '([{':
                depth += 1
            elif c in ')]}':
                depth -= 1
            i += 1
        cont = line.rstrip().endswith('\\') and not triple
        if triple or depth > 0 or cont:
            continue
        out.append('\n'.join(buf))
        buf, depth = [], 0
    if buf:
        out.append('\n'.join(buf))
    return out


def _fix_print(logical):
    first, _, tail = logical.partition('\n')
    m = _PRINT_RE.match(first)
    if not m:
        return logical
    rest = m.group('rest') + ('\n' + tail if tail else '')
    if rest.lstrip().startswith('(') and not m.group('chev'):
        # already a call-looking form: `print (a)` / print('x' 'y') — valid in both dialects
        probe = rest.strip()
        depth = 0
        for k, c in enumerate(probe):
            depth += c == '('
            depth -= c == ')'
            if depth == 0:
                if probe[k + 1:].strip() == '' or probe[k + 1:].lstrip().startswith('#'):
                    return logical
                break
    body = rest.strip()
    comment = ''
    last_nl = body.rfind('\n') + 1
    in_str, k = None, last_nl
    if '\n' not in body:                      # a trailing comment on a one-line print moves behind the ')'
        while k < len(body):
            c = body[k]
            if in_str:
                if c == '\\':
                    k += 1
                elif c == in_str:
                    in_str = None
            elif c in '\'"':
                in_str = c
            elif c == '#':
                body, comment = body[:k].rstrip(), ' ' + body[k:]
                break
            k += 1
    end = ''
    if body.endswith(','):
        body, end = body[:-1], ", end=' '"
    target = ''
    if m.group('chev'):
        target = ', file=' + m.group('chev').strip()[2:-1].strip()
    return '%sprint(%s%s%s)%s' % (m.group('head'), body, end, target, comment)


_TEXT_SUBS = [
    (re.compile(r'\bexcept\s+([\w.]+|\([^)]*\))\s*,\s*(\w+)\s*:'), r'except \1 as \2:'),              # T2
    (re.compile(r'\blambda\s*\(\s*(\w+)\s*\)\s*:'), r'lambda \1:'),                                    # T3
    (re.compile(r'\bexec\s+(?!\()(.+?)\s+in\s+(\w+)\s*$', re.S | re.M), r'exec(\1, \2)'),              # T4
    (re.compile(r'^from Queue import', re.M), 'from queue import'),                                     # T5
    (re.compile(r'^from StringIO import StringIO', re.M), 'from io import StringIO'),
    (re.compile(r'^(\s*)import cPickle as pickle', re.M), r'\1import pickle'),
    (re.compile(r'^from collections import ((?:Sequence|Sized|Set|Mapping|Iterable|Hashable)(?:\s*,\s*'
                r'(?:Sequence|Sized|Set|Mapping|Iterable|Hashable))*)\s*$', re.M), r'from collections.abc import \1'),
    (re.compile(r'^from multiprocessing\.queues import Queue', re.M), 'from _refshim_compat import Queue'),
    (re.compile(r'\.iteritems\(\)'), '.items()'),                                                       # T6
    (re.compile(r'\.itervalues\(\)'), '.values()'),
    (re.compile(r'\.iterkeys\(\)'), '.keys()'),
    (re.compile(r'\.next\(\)'), '.__next__()'),
    (re.compile(r'^from array import array', re.M), 'from _refshim_compat import array'),               # T7
    (re.compile(r"open\(filename, 'rb'\)"), "open(filename, 'r')"),                                     # T8
]


def py2_text_to_py3(src):
    src = '\n'.join(_fix_print(l) for l in _logical_lines(src))
    for rx, rep in _TEXT_SUBS:
        src = rx.sub(rep, src)
    return src


# --------------------------------------------------------------------------- AST pass
class _Py2Semantics(ast.NodeTransformer):
    def visit_BinOp(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):                                                                # A1
            return ast.copy_location(
                ast.Call(func=ast.Name(id='__py2div__', ctx=ast.Load()), args=[node.left, node.right], keywords=[]),
                node)
        return node

    def visit_AugAssign(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):                                                                # A1
            import copy
            load = copy.deepcopy(node.target)
            for n in ast.walk(load):
                if hasattr(n, 'ctx'):
                    n.ctx = ast.Load()
            call = ast.Call(func=ast.Name(id='__py2idiv__', ctx=ast.Load()), args=[load, node.value], keywords=[])
            return ast.copy_location(ast.Assign(targets=[node.target], value=call), node)
        return node

    def visit_ClassDef(self, node):
        self.generic_visit(node)
        extra = []
        for stmt in list(node.body):
            if isinstance(stmt, ast.FunctionDef) and stmt.name == '__nonzero__':                        # A2
                extra.append(ast.Assign(targets=[ast.Name(id='__bool__', ctx=ast.Store())],
                                        value=ast.Name(id='__nonzero__', ctx=ast.Load())))
            if isinstance(stmt, ast.FunctionDef) and stmt.name in ('__div__', '__idiv__', '__rdiv__'):   # A5
                extra.append(ast.Assign(targets=[ast.Name(id=stmt.name.replace('div', 'truediv'), ctx=ast.Store())],
                                        value=ast.Name(id=stmt.name, ctx=ast.Load())))
            if (isinstance(stmt, ast.Assign) and len(stmt.targets) == 1 and isinstance(stmt.targets[0], ast.Name)
                    and stmt.targets[0].id == '__metaclass__'):                                         # A3
                node.keywords.append(ast.keyword(arg='metaclass', value=stmt.value))
                node.body.remove(stmt)
        node.body.extend(extra)
        if not node.body:
            node.body.append(ast.Pass())
        return node

    def visit_Call(self, node):
        self.generic_visit(node)
        f = node.func
        if (isinstance(f, ast.Attribute) and f.attr in ('keys', 'values', 'items')
                and not node.args and not node.keywords):                                               # A4
            return ast.copy_location(
                ast.Call(func=ast.Name(id='__py2list__', ctx=ast.Load()), args=[node], keywords=[]), node)
        return node


def _py2div(a, b):
    """Python-2 `/`: floor division when both operands are integers (bool and numpy integer scalars
    included), true division otherwise.  Integer numpy ARRAYS would also floor-divide under Python 2;
    the path never divides one (its arrays are float / complex), so that case is not emulated."""
    import numbers
    import numpy as np
    ints = (numbers.Integral, np.integer)
    if isinstance(a, ints) and isinstance(b, ints):
        return a // b
    return a / b


def _py2idiv(a, b):
    """Python-2 `a /= b`: the object's own `__idiv__` (in place) when it has one, else `a = a / b`."""
    own = getattr(type(a), '__idiv__', None)
    if own is not None:
        return own(a, b)
    return _py2div(a, b)


def _py2list(x):
    # dict views become lists; anything else that happens to have a .keys()/.values()/.items() method and already
    # returns a list/dict/proxy is passed through untouched
    if isinstance(x, (type({}.keys()), type({}.values()), type({}.items()))):
        return list(x)
    return x


def _listing(fn):
    @functools.wraps(fn)
    def wrapper(*a, **k):
        return list(fn(*a, **k))
    return wrapper


def _py2_map(fn, *seqs):
    if fn is None:
        return list(zip(*seqs)) if len(seqs) > 1 else list(seqs[0])
    return list(builtins.map(fn, *seqs))


_NAMES = {
    '__py2div__': _py2div, '__py2idiv__': _py2idiv, '__py2list__': _py2list,
    'map': _py2_map, 'filter': _listing(builtins.filter), 'zip': _listing(builtins.zip),
    'range': _listing(builtins.range), 'xrange': builtins.range,
    'basestring': str, 'unicode': str, 'long': int, 'reduce': functools.reduce, 'raw_input': input,
}


def compile_py2(src, filename):
    tree = ast.parse(py2_text_to_py3(src), filename)
    tree = ast.fix_missing_locations(_Py2Semantics().visit(tree))
    return compile(tree, filename, 'exec')


# --------------------------------------------------------------------------- compat + Biopython stand-ins
def _install_compat():
    if '_refshim_compat' in sys.modules:
        return
    import array as _array
    compat = types.ModuleType('_refshim_compat')

    class _CharArray(list):
        '''py2 array('c', s): a mutable string (DP/SecStructures.py:367)'''
        def tostring(self): return ''.join(self)

    def array(typecode, init=None):
        if typecode == 'c':
            return _CharArray(init or '')
        if init is None:
            return _array.array(typecode)
        if isinstance(init, str):
            init = init.encode('latin-1')
        return _array.array(typecode, init)
    compat.array = array

    import multiprocessing
    import multiprocessing.queues

    class Queue(multiprocessing.queues.Queue):
        '''py2 signature Queue(maxsize=0); 3.x wants the context as a keyword'''
        def __init__(self, maxsize=0):
            multiprocessing.queues.Queue.__init__(self, maxsize, ctx=multiprocessing.get_context('fork'))
    compat.Queue = Queue
    sys.modules['_refshim_compat'] = compat


_COMPLEMENT = str.maketrans('ACGTRYSWKMBDHVNacgtryswkmbdhvn', 'TGCAYRSWMKVHDBNtgcayrswmkvhdbn')


class Seq(object):
    """Stand-in for Bio.Seq.Seq (Biopython 1.66): an immutable string with an alphabet."""
    def __init__(self, data='', alphabet=None):
        self._data = str(data)
        self.alphabet = alphabet

    def __str__(self): return self._data
    def __repr__(self): return 'Seq(%r)' % self._data
    def __len__(self): return len(self._data)
    def __hash__(self): return hash(self._data)
    def __eq__(self, other): return self._data == str(other)
    def __ne__(self, other): return self._data != str(other)
    def __iter__(self): return iter(self._data)
    def __contains__(self, x): return str(x) in self._data
    def __add__(self, other): return Seq(self._data + str(other), self.alphabet)
    def __radd__(self, other): return Seq(str(other) + self._data, self.alphabet)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return Seq(self._data[i], self.alphabet)
        return self._data[i]

    def upper(self): return Seq(self._data.upper(), self.alphabet)
    def lower(self): return Seq(self._data.lower(), self.alphabet)
    def count(self, sub, *a): return self._data.count(str(sub), *a)
    def find(self, sub, *a): return self._data.find(str(sub), *a)
    def complement(self): return Seq(self._data.translate(_COMPLEMENT), self.alphabet)
    def reverse_complement(self): return Seq(self._data.translate(_COMPLEMENT)[::-1], self.alphabet)
    def tostring(self): return self._data


class SeqRecord(object):
    """Stand-in for Bio.SeqRecord.SeqRecord."""
    def __init__(self, seq, id='<unknown id>', name='<unknown name>', description='<unknown description>',
                 dbxrefs=None, features=None, annotations=None, letter_annotations=None):
        self.seq = seq
        self.id = id
        self.name = name
        self.description = description
        self.dbxrefs = dbxrefs or []
        self.features = features or []
        self.annotations = annotations or {}
        self.letter_annotations = letter_annotations or {}

    def __len__(self): return len(self.seq)
    def __str__(self): return str(self.seq)
    def __iter__(self): return iter(self.seq)
    def __bool__(self): return True

    def __getitem__(self, i):
        if isinstance(i, slice):
            return SeqRecord(self.seq[i], id=self.id, name=self.name, description=self.description)
        return self.seq[i]

    def upper(self):
        return SeqRecord(self.seq.upper(), id=self.id, name=self.name, description=self.description,
                         annotations=dict(self.annotations))

    def reverse_complement(self, id=False, name=False, description=False, **_):
        return SeqRecord(self.seq.reverse_complement(), id=self.id if id is True else (id or '<unknown id>'),
                         name=self.name if name is True else (name or '<unknown name>'),
                         description=self.description if description is True else (description or ''))


def _fasta_parse(handle, fmt='fasta', alphabet=None):
    assert fmt == 'fasta', 'refshim Bio.SeqIO stand-in reads FASTA only'
    close = False
    if isinstance(handle, str):
        handle, close = open(handle), True
    try:
        title, chunks = None, []
        for line in handle:
            line = line.rstrip('\r\n')
            if line.startswith('>'):
                if title is not None:
                    yield _fasta_record(title, chunks, alphabet)
                title, chunks = line[1:], []
            elif title is not None:
                chunks.append(line.strip())
        if title is not None:
            yield _fasta_record(title, chunks, alphabet)
    finally:
        if close:
            handle.close()


def _fasta_record(title, chunks, alphabet):
    rid = title.split(None, 1)[0] if title.strip() else ''
    return SeqRecord(Seq(''.join(chunks), alphabet), id=rid, name=rid, description=title)


def _install_bio():
    if 'Bio' in sys.modules:
        return
    bio = types.ModuleType('Bio')
    bio.__path__ = []
    m_seq = types.ModuleType('Bio.Seq')
    m_seq.Seq = Seq
    m_rec = types.ModuleType('Bio.SeqRecord')
    m_rec.SeqRecord = SeqRecord
    m_alpha = types.ModuleType('Bio.Alphabet')
    m_alpha.__path__ = []

    class _Alphabet(object):
        def __init__(self, letters): self.letters = letters
    m_iupac = types.ModuleType('Bio.Alphabet.IUPAC')
    m_iupac.unambiguous_dna = _Alphabet('GATC')
    m_iupac.ambiguous_dna = _Alphabet('GATCRYWSMKHBVDN')
    m_alpha.IUPAC = m_iupac
    for n in ('generic_alphabet', 'generic_dna', 'generic_protein'):
        setattr(m_alpha, n, _Alphabet(None))
    for n in ('ProteinAlphabet', 'DNAAlphabet', 'NucleotideAlphabet'):
        setattr(m_alpha, n, _Alphabet)
    m_io = types.ModuleType('Bio.SeqIO')
    m_io.parse = _fasta_parse
    bio.Seq, bio.SeqRecord, bio.Alphabet, bio.SeqIO = m_seq, m_rec, m_alpha, m_io
    sys.modules.update({'Bio': bio, 'Bio.Seq': m_seq, 'Bio.SeqRecord': m_rec, 'Bio.Alphabet': m_alpha,
                        'Bio.Alphabet.IUPAC': m_iupac, 'Bio.SeqIO': m_io})


# --------------------------------------------------------------------------- importer
class _RefLoader(importlib.abc.Loader):
    def __init__(self, path, is_pkg):
        self.path, self.is_pkg = path, is_pkg

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        with open(self.path, 'rb') as f:
            raw = f.read()
        src = raw.decode('utf-8')
        src = re.sub(r'^#.*coding[=:].*$', '#', src, count=1, flags=re.M)
        here = os.path.dirname(self.path)

        def sibling(name):
            return os.path.isfile(os.path.join(here, name + '.py')) or \
                os.path.isfile(os.path.join(here, name, '__init__.py'))
        # (T9) py2 implicit relative imports: a sibling module wins over a top-level one
        src = re.sub(r'^(\s*)from (\w+) import', lambda m: '%sfrom .%s import' % (m.group(1), m.group(2))
                     if sibling(m.group(2)) else m.group(0), src, flags=re.M)
        src = re.sub(r'^(\s*)import (\w+)\s*$', lambda m: '%sfrom . import %s' % (m.group(1), m.group(2))
                     if sibling(m.group(2)) else m.group(0), src, flags=re.M)
        module.__dict__.update(_NAMES)
        module.__file__ = self.path
        exec(compile_py2(src, self.path), module.__dict__)


class _RefFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        top = fullname.split('.')[0]
        if top not in _ROOTS:
            return None
        rel = fullname.split('.')[1:]
        base = os.path.join(_ROOTS[top], *rel)
        if os.path.isdir(base) and os.path.isfile(os.path.join(base, '__init__.py')):
            spec = importlib.util.spec_from_loader(fullname, _RefLoader(os.path.join(base, '__init__.py'), True),
                                                   is_package=True)
            spec.submodule_search_locations = [base]
            return spec
        if os.path.isfile(base + '.py'):
            return importlib.util.spec_from_loader(fullname, _RefLoader(base + '.py', False))
        return None


_installed = False


def install():
    """Make `import DegenPrimer.SearchEngine` etc. resolve to the reference's own sources."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError('reference sources not found under %s (refshim only runs in the build container)'
                           % REFERENCE_ROOT)
    _install_compat()
    _install_bio()
    sys.meta_path.insert(0, _RefFinder())
    _installed = True


def load():
    """Import the hot-path modules; returns a namespace with the reference's modules and the stand-in types."""
    install()
    cwd = os.getcwd()
    os.chdir(os.path.join(REFERENCE_ROOT, 'data'))      # DP/UnifiedNN.py:103-110 looks for the tables in './' first
    try:
        import DegenPrimer.TD_Functions as tdf
        import DegenPrimer.SecStructures as sec
        import DegenPrimer.SearchEngine as se
        import DegenPrimer.Primer as primer
        import DegenPrimer.WorkCounter as wc
    finally:
        os.chdir(cwd)
    return types.SimpleNamespace(tdf=tdf, SecStructures=sec, SearchEngine=se, Primer=primer, WorkCounter=wc,
                                 Seq=Seq, SeqRecord=SeqRecord)
