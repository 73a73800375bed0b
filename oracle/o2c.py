"""ORACLE (test infrastructure) — ctypes wrapper of oracle/o2_scan.c."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, 'o2_scan.c')
LIB = os.path.join(_HERE, '_build', 'libo2scan.so')


def build(force=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(['gcc', '-O3', '-fopenmp', '-shared', '-fPIC', SRC, '-o', LIB])
    return LIB


_lib = None


def find(fwd_seq, master, mismatches):
    """Same convention as search_o2.find: (fwd_matches, rev_matches) int64 arrays."""
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.o2_scan.restype = C.c_int
        _lib.o2_scan.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_int, C.c_int,
                                 C.c_void_p, C.c_int64, C.POINTER(C.c_int64),
                                 C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
    t = fwd_seq if isinstance(fwd_seq, bytes) else fwd_seq.encode('ascii')
    cap = 1 << 16
    while True:
        f = np.zeros(cap, dtype=np.int64)
        r = np.zeros(cap, dtype=np.int64)
        nf, nr = C.c_int64(), C.c_int64()
        rc = _lib.o2_scan(t, len(t), master.encode('ascii'), len(master), int(mismatches),
                          f.ctypes.data, cap, C.byref(nf), r.ctypes.data, cap, C.byref(nr))
        if rc:
            raise ValueError('o2_scan failed with %d' % rc)
        if nf.value <= cap and nr.value <= cap:
            return f[:nf.value].copy(), r[:nr.value].copy()
        cap = max(nf.value, nr.value)
