"""dp_variant_tm_static — the entry point that allocates nothing (no counterpart in the reference):
per-variant Tm with its staging memory in a file-scope static __device__
array.  Called through the C-ABI, checked against dp_variant_tm on two calls in a row (the second one smaller,
so stale contents of the shared array would show), plus the size limit and the error path.  This file sorts
last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

COND = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0, DMSO=2.0)


def _both(ctx, cond, seqs, conc=1.5e-6, cdeg=0.5):
    """(rc, Tm by dp_variant_tm_static, Tm by dp_variant_tm) for equal-length sequences."""
    n, L = len(seqs), len(seqs[0])
    blob = ''.join(seqs).encode('ascii')
    ptr = C.cast(C.c_char_p(blob), C.c_void_p)
    got = np.full(n, np.nan)
    rc = ctx.lib.dp_variant_tm_static(ctx._h, n, L, ptr, conc, cdeg, C.byref(cond), got.ctypes.data_as(C.c_void_p))
    want = np.full(n, np.nan)
    ctx._check(ctx.lib.dp_variant_tm(ctx._h, n, L, ptr, conc, cdeg, C.byref(cond), want.ctypes.data_as(C.c_void_p)))
    return rc, got, want


def test_static_entry_point_equals_dp_variant_tm():
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import _native as nat
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    tdf.PCR_P.set(COND)
    ctx = tdf._ctx()
    cond = tdf.conditions()
    max_seqs, max_bytes, scratch = C.c_uint32(), C.c_uint64(), C.c_uint64()
    assert ctx.lib.dp_variant_tm_static_limits(C.byref(max_seqs), C.byref(max_bytes), C.byref(scratch)) == 0
    assert (max_seqs.value, max_bytes.value) == (4096, 262144) and scratch.value <= 512 * 1024
    # call 1: 4096 variants of a 26-mer (lower case too: both paths fold case)
    first = unambiguous_sequences('AGRGTTYGATYMTGGYTYWGKSRYMA')
    assert len(first) == 4096                     # exactly the limit
    first[5] = first[5].lower()
    rc, got, want = _both(ctx, cond, first)
    assert rc == 0 and np.isfinite(want).all() and want.std() > 0.5
    assert np.abs(got - want).max() < 1e-9 and np.array_equal(got, want)       # the tests' tolerance; in fact bit-identical
    # call 2, straight after it: fewer and shorter sequences over the same array
    second = unambiguous_sequences('KACGGYTAMCTTRTTAYGASTW')
    rc, got2, want2 = _both(ctx, cond, second, conc=2.5e-7, cdeg=0.3)
    assert rc == 0 and len(got2) == 128 and np.abs(got2 - want2).max() < 1e-9 and np.array_equal(got2, want2)
    # and the first call again gives the first answer again
    rc, got3, _ = _both(ctx, cond, first)
    assert rc == 0 and np.array_equal(got3, got)
    # beyond the static array: refused with a message that names the limit, nothing is computed
    big = ['ACGT' * 16] * 4097
    rc, out, _ = _both(ctx, cond, big[:1])
    assert rc == 0
    blob = ''.join(big).encode()
    tm = np.zeros(len(big))
    rc = ctx.lib.dp_variant_tm_static(ctx._h, len(big), 64, C.cast(C.c_char_p(blob), C.c_void_p), 1e-6, 0.5,
                                      C.byref(cond), tm.ctypes.data_as(C.c_void_p))
    msg = ctx.lib.dp_last_error(ctx._h).decode()
    assert rc == nat.DP_ERR_UNSUPPORTED and '4096' in msg and '262144' in msg and not tm.any()
    rc = ctx.lib.dp_variant_tm_static(ctx._h, 4096, 65, C.cast(C.c_char_p(blob + b'A' * 4096), C.c_void_p), 1e-6, 0.5,
                                      C.byref(cond), tm.ctypes.data_as(C.c_void_p))
    assert rc == nat.DP_ERR_UNSUPPORTED                                          # 4096 x 65 bytes > 262144
    # a letter that is not an unambiguous base: same error class as dp_variant_tm, position reported;
    # the status words are cleared by the next call
    bad = list(second[:8])
    bad[3] = bad[3][:7] + 'N' + bad[3][8:]
    blob = ''.join(bad).encode()
    tm = np.zeros(8)
    rc = ctx.lib.dp_variant_tm_static(ctx._h, 8, 22, C.cast(C.c_char_p(blob), C.c_void_p), 1e-6, 0.5, C.byref(cond),
                                      tm.ctypes.data_as(C.c_void_p))
    assert rc == nat.DP_ERR_ALPHABET and 'byte %d' % (3 * 22 + 7) in ctx.lib.dp_last_error(ctx._h).decode()
    assert ctx.lib.dp_variant_tm(ctx._h, 8, 22, C.cast(C.c_char_p(blob), C.c_void_p), 1e-6, 0.5, C.byref(cond),
                                 tm.ctypes.data_as(C.c_void_p)) == nat.DP_ERR_ALPHABET
    rc, got4, want4 = _both(ctx, cond, second)
    assert rc == 0 and np.array_equal(got4, want4)
