"""CPU tests of the host side: the C-ABI library loads and exports every symbol
the header declares, the host mirrors behave like the reference's classes, and
the product never routes through the oracle."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, 'include', 'degenprimer_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(dp_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_declared_symbol(native_lib):
    lib = ctypes.CDLL(native_lib)
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), 'missing export %s' % n
    from degenprimer_b200 import _native
    assert sorted(_native.EXPORTS) == names
    h = _native.load()
    assert h.dp_abi_version() == 1


def test_no_gpu_means_loud_failure(native_lib):
    """No compute call without a GPU: dp_create must fail with a message, never fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from degenprimer_b200 import _native
    with pytest.raises(_native.NativeError) as e:
        _native.Context(0)
    assert 'no CPU path' in str(e.value) or 'CUDA' in str(e.value)


def test_missing_library_is_loud(tmp_path):
    code = ('import sys; sys.path.insert(0, %r)\n'
            'import degenprimer_b200._native as n\n'
            'import degenprimer_b200.build as b\n'
            'n.LIB_PATH = %r; b._stale = lambda: False; n._lib = None\n'
            'try:\n    n.load()\nexcept RuntimeError as e:\n    print("RAISED", e)\n' % (ROOT, str(tmp_path / 'nope.so')))
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True).stdout
    assert 'RAISED' in out and 'no CPU fallback' in out


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'degenprimer_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f
                assert 'oracle.' not in text.replace('oracle/', ''), f


def test_host_mirrors_match_the_oracle_on_host_logic():
    from degenprimer_b200.SeqUtils import unambiguous_sequences, reverse_complement, Seq, SeqRecord, pretty_rec_name
    from degenprimer_b200.SecStructures import Dimer
    from degenprimer_b200.Primer import Primer
    from oracle import sequtils as osu
    from oracle.structures import Dimer as ODimer
    from oracle.primer import Primer as OPrimer, SeqRecord as OSeqRecord
    rng = np.random.default_rng(0)
    for _ in range(50):
        s = ''.join(rng.choice(list('ACGTRYSWKMBDHVN'), int(rng.integers(1, 9))))
        assert unambiguous_sequences(s) == osu.unambiguous_sequences(s)
        assert reverse_complement(s) == osu.reverse_complement(s)
        m = sorted(set(rng.integers(0, 30, size=int(rng.integers(1, 25))).tolist()))
        assert [d.fwd_matches for d in Dimer(m).alternatives] == [d.fwd_matches for d in ODimer(m).alternatives]
    with pytest.raises(KeyError):
        reverse_complement('ACGTX')
    a = Primer(SeqRecord(Seq('GAASGCRAAKATYGGGAAC'), id='p'), 2e-6)
    a.generate_components()
    b = OPrimer(OSeqRecord('GAASGCRAAKATYGGGAAC', id='p'), 2e-6)
    assert [r.id for r in a.seq_records] == [r.id for r in b.seq_records]
    assert a.str_sequences == b.str_sequences and a.concentration == b.concentration
    assert len(a) == 19 and a.num_components == 16 and 'GAAGGCAAAGATCGGGAAC' in a
    assert pretty_rec_name(SeqRecord(Seq('A'), id='x', description='x chr=1/2')) == 'chr-1-2'
    d = Dimer.from_sequences('ACGTACGT', 'ACGAACGT')
    assert d.fwd_matches == (0, 1, 2, 4, 5, 6, 7) and d.fwd_max == 7 and d.rev_min == 0


def test_tables_and_conditions_match_the_oracle():
    from degenprimer_b200 import TD_Functions as tdf, _native as nat
    from degenprimer_b200.UnifiedNN import UnifiedNN
    from oracle import unified_nn as onn, conditions as ocond
    t = UnifiedNN.device_tables()
    comp = dict(zip('ACGT', 'TGCA'))
    for xi, x in enumerate(nat.BASES):
        for yi, y in enumerate(nat.BASES):
            for wi, w in enumerate(nat.LETTERS):
                i = (xi * 4 + yi) * 15 + wi
                p = onn.int_dPar_37(x + y, comp[x] + w)
                assert (t.stack_dH[i], t.stack_dS[i]) == (p.dH, p.dS)
                q = onn.int_dPar_37(comp[y] + w, y + x)
                assert abs(t.rstack_dH[i] - q.dH) < 1e-12 and abs(t.rstack_dS[i] - q.dS) < 1e-12
    for n in range(3, 40):
        assert t.loop_dS[n] == onn.internal_loop_dS_37('A' * (n + 2))
    c = dict(Na=40e-3, Mg=2e-3, dNTP=0.2e-3, DNA=2e-9, DMSO=1.0, PCR_T=57.0)
    tdf.PCR_P.set(c)
    ocond.PCR_P.set(c)
    assert tdf.PCR_P.Na_eq_log == ocond.PCR_P.Na_eq_log
    cut = tdf.max_dG_for_min_K(57.0)
    assert tdf.equilibrium_constant(cut, 57.0) >= 100 > tdf.equilibrium_constant(np.nextafter(cut, np.inf), 57.0)
    with tdf.AcquireParameters():
        assert tdf.PCR_P is not tdf._PCR_P and tdf.PCR_P.PCR_T == 57.0
    with pytest.raises(ValueError):
        tdf.PCR_P.set(dict(Mg=1e-4, dNTP=1e-3))
    tdf.PCR_P.set(dict(Mg=1.5e-3, dNTP=0.1e-3))
    ref_dir = '/root/reference/data'
    if os.path.isdir(ref_dir):            # built-in numbers == the reference's csv files
        before = {k: (v.dH, v.dS) for k, v in UnifiedNN.internal_NN.items() if len(k) == 5 and k[2] == '/'}
        UnifiedNN.load_tables(ref_dir)
        after = {k: (v.dH, v.dS) for k, v in UnifiedNN.internal_NN.items() if len(k) == 5 and k[2] == '/'}
        UnifiedNN.load_tables()
        assert all(before[k] == after[k] for k in after)


def test_work_counter_and_shelf(tmp_path):
    from degenprimer_b200.WorkCounter import WorkCounter
    from degenprimer_b200.tmpStorage import to_shelf, from_shelf, cleanup_file
    c = WorkCounter()
    c.set_subwork(2, (10, 30))
    c[0].set_work(5)
    c[0].count(5)
    assert abs(c.percent() - 25.0) < 1e-9
    c[1].set_work(4)
    c[1].done()
    assert abs(c.percent() - 100.0) < 1e-9
    assert to_shelf(None) is None
    path = to_shelf(([(5, ('a',))], None))
    assert from_shelf(path) == ([(5, ('a',))], None)
    cleanup_file(path)
    assert not os.path.exists(path)


def test_search_engine_surface():
    from degenprimer_b200 import SearchEngine as se
    import inspect
    eng = se.SearchEngine()
    for name in ('find_matches', 'compile_duplexes', 'compile_duplexes_mp', 'find', 'batch_find',
                 'mp_better', 'set_max_chunk', '_check_length_inequality'):
        assert hasattr(eng, name)
    assert list(inspect.signature(eng.find_matches).parameters) == ['counter', 'template', 'primer', 'mismatches']
    assert list(inspect.signature(eng.compile_duplexes_mp).parameters) == \
        ['counter', 'fwd_seq', 'rev_seq', 'primer', 't_len', 'p_len', 'fwd_matches', 'rev_matches', 'num_jobs']
    assert list(inspect.signature(se.find).parameters) == ['counter', 'template', 'primer', 'mismatches', 'abort_event']
    assert se.mp_better(10) is False
    with pytest.raises(ValueError):
        se.SearchEngine._check_length_inequality(5, 10)
    with pytest.raises(ValueError):
        se.SearchEngine._check_length_inequality(5, 0)


def test_specialised_scan_kernel_compiles_without_a_device(native_lib, tmp_path):
    """dp_jit_selftest: source generation + NVRTC compile of the run-time specialised scan
    kernel (sm_100a) for a plain pair, a degenerate pair and a 20-plex; the source has one
    group-count prefilter per (primer, strand) pattern, feeding the warp's candidate queue."""
    lib = ctypes.CDLL(native_lib)
    lib.dp_last_error.restype = ctypes.c_char_p
    from degenprimer_b200.synthetic import TP53_F, TP53_R, P27F_D128, P1492R_D128, multiplex_primers
    for masters, k, units in (([TP53_F, TP53_R], 1, 1), ([P27F_D128, P1492R_D128], 3, 1), (multiplex_primers(), 3, 1)):
        arr = (ctypes.c_char_p * len(masters))(*[m.encode() for m in masters])
        ks = (ctypes.c_int32 * len(masters))(*([k] * len(masters)))
        src = str(tmp_path / 'unit.cu')
        rc = lib.dp_jit_selftest(len(masters), arr, ks, 3, src.encode(), None)
        if rc < 0 and b'libnvrtc' in lib.dp_last_error(None):
            pytest.skip('NVRTC is not installed here')
        assert rc == units, lib.dp_last_error(None)
        text = open(src).read()
        assert text.count('qn = dp_queue_push(A, s_pat') == 2 * len(masters)   # one prefilter per (primer, strand)
        assert text.count('//PLAN pattern') == 2 * len(masters)
        assert ('shared-planes' in text) == (len(masters) == 20)               # a multiplex keeps its match planes in registers
    bad = (ctypes.c_char_p * 1)(b'ACGTZ')
    assert lib.dp_jit_selftest(1, bad, (ctypes.c_int32 * 1)(0), 3, None, None) < 0
