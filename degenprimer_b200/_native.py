"""ctypes binding of include/degenprimer_b200.h.

There is no fallback: if libdegenprimer_b200.so is missing, or no sm_100 GPU is
usable, every entry point raises.  The library is loaded lazily so importing
the package (and forking, as the reference's Manager processes do —
BioUtils/Tools/UMP.py:83-91) never creates a CUDA context by itself.
"""
import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libdegenprimer_b200.so')
_LIB_OVERRIDE = os.environ.get('DEGENPRIMER_B200_LIB')      # tuning: a library built with other kernel parameters

DP_OK = 0
DP_ERR_CUDA, DP_ERR_ARG, DP_ERR_ALPHABET, DP_ERR_LENGTH, DP_ERR_UNSUPPORTED = -1, -2, -3, -4, -5
DP_ERR_IO, DP_ERR_NOMEM, DP_ERR_ABORTED, DP_ERR_STATE = -6, -7, -8, -9
DP_MAX_PRIMER_LEN = 64
DP_MAX_PRIMERS = 64
DP_STRAND_FWD, DP_STRAND_REV, DP_STRAND_BOTH = 1, 2, 3
DP_ALGO_AUTO, DP_ALGO_POPC, DP_ALGO_BITSLICE, DP_ALGO_JIT = 0, 1, 2, 3

LETTERS = 'ACTGRYSWKMBDHVN'      # letter index order of the ABI
BASES = 'ACTG'                   # 2-bit plane code order


class NativeError(RuntimeError):
    def __init__(self, code, message):
        RuntimeError.__init__(self, 'degenprimer_b200 native error %d: %s' % (code, message))
        self.code = code
        self.message = message


class Aborted(Exception):
    """The abort flag was raised; the facade turns this into `return None`."""


class dp_template_info(C.Structure):
    _fields_ = [('n_records', C.c_uint32), ('reserved', C.c_uint32),
                ('total_bases', C.c_uint64), ('padded_bases', C.c_uint64),
                ('n_exception_runs', C.c_uint64), ('device_bytes', C.c_uint64)]


class _Stats(C.Structure):
    """ctypes struct that also reads like the dict the callers expect."""

    def __getitem__(self, name):
        return getattr(self, name)

    def keys(self):
        return [f for f, _ in self._fields_]

    def as_dict(self):
        return {f: getattr(self, f) for f, _ in self._fields_}

    def __repr__(self):
        return repr(self.as_dict())


class dp_scan_stats(_Stats):
    _fields_ = [('windows', C.c_uint64), ('candidates', C.c_uint64), ('hits', C.c_uint64),
                ('ms_scan', C.c_float), ('ms_total', C.c_float),
                ('ms_main', C.c_float), ('reserved', C.c_float),
                ('launches', C.c_uint32), ('algo', C.c_uint32)]


class dp_dg_stats(_Stats):
    _fields_ = [('duplexes', C.c_uint64), ('dimers_tried', C.c_uint64), ('dimers_kept', C.c_uint64),
                ('ms_kernel', C.c_float), ('ms_total', C.c_float),
                ('launches', C.c_uint32), ('reserved', C.c_uint32)]


class dp_nn_tables(C.Structure):
    _fields_ = [('stack_dH', C.c_double * (4 * 4 * 15)), ('stack_dS', C.c_double * (4 * 4 * 15)),
                ('rstack_dH', C.c_double * (4 * 4 * 15)), ('rstack_dS', C.c_double * (4 * 4 * 15)),
                ('dangling_dH', C.c_double * (2 * 4 * 4 * 4)), ('dangling_dS', C.c_double * (2 * 4 * 4 * 4)),
                ('loop_dS', C.c_double * 64),
                ('ini_dH', C.c_double), ('ini_dS', C.c_double), ('ter_dH', C.c_double), ('ter_dS', C.c_double),
                ('terminal_mismatch', C.c_double), ('dS_Na', C.c_double), ('R', C.c_double),
                ('K0', C.c_double), ('T_DMSO', C.c_double)]


class dp_hairpin_tables(C.Structure):
    _fields_ = [('loop_dG', C.c_double * 64), ('tri_dH', C.c_double * 1024), ('tri_dG', C.c_double * 1024),
                ('tetra_dH', C.c_double * 4096), ('tetra_dG', C.c_double * 4096), ('iK37', C.c_double)]


class dp_conditions(C.Structure):
    _fields_ = [('PCR_T', C.c_double), ('Na_eq_log', C.c_double), ('max_dimer_dG', C.c_double),
                ('max_dG_for_K', C.c_double), ('DNA', C.c_double), ('DMSO', C.c_double)]


HIT_DTYPE = np.dtype([('pos', '<u8'), ('record', '<u4'), ('primer', '<u2'),
                      ('strand', 'u1'), ('score3', 'u1')], align=True)
DIMER_DTYPE = np.dtype([('hit', '<u4'), ('variant', '<u4'), ('match_mask', '<u8'), ('dG', '<f8')],
                       align=True)
WIRE_DTYPE = np.dtype([('g', '<u8'), ('record', '<u4'), ('pattern', '<u2'), ('score3', '<u2'),
                       ('pair_base', '<u4'), ('reserved', '<u4')], align=True)
WIRE_HEADER_BYTES = 64
BEST_STRUCTURE_DTYPE = np.dtype([('best_mask', '<u8'), ('best_dG', '<f8')], align=True)
FRONT_DIMER_DTYPE = np.dtype([('full_mask', '<u8'), ('best_mask', '<u8'), ('best_dG', '<f8')], align=True)
assert HIT_DTYPE.itemsize == 16 and DIMER_DTYPE.itemsize == 24 and WIRE_DTYPE.itemsize == 24
assert FRONT_DIMER_DTYPE.itemsize == 24
PAIR_DTYPE = np.dtype([('fwd', '<u4'), ('rev', '<u4')])

_lib = None
_lock = threading.Lock()


def _declare(lib):
    vp, cp, u32, u64, i32 = C.c_void_p, C.c_char_p, C.c_uint32, C.c_uint64, C.c_int
    P = C.POINTER
    sig = {
        'dp_abi_version': (i32, []),
        'dp_device_count': (i32, [P(i32)]),
        'dp_create': (vp, [i32]),
        'dp_create_error': (cp, []),
        'dp_destroy': (None, [vp]),
        'dp_last_error': (cp, [vp]),
        'dp_free': (None, [vp]),
        'dp_abort_requested': (i32, [vp, i32]),
        'dp_device_info': (i32, [vp, C.c_char_p, P(i32), P(i32), P(i32), P(i32), P(u64)]),
        'dp_template_from_records': (i32, [vp, u32, P(C.c_void_p), P(u64), P(vp)]),
        'dp_template_from_ascii': (i32, [vp, C.c_void_p, u64, P(vp)]),
        'dp_template_from_fasta': (i32, [vp, cp, P(vp)]),
        'dp_template_set_record_name': (i32, [vp, u32, cp, cp]),
        'dp_template_save': (i32, [vp, vp, cp]),
        'dp_template_load': (i32, [vp, cp, P(vp)]),
        'dp_template_image_size': (i32, [vp, P(u64)]),
        'dp_template_save_mem': (i32, [vp, vp, C.c_void_p, u64]),
        'dp_template_load_mem': (i32, [vp, C.c_void_p, u64, i32, P(vp)]),
        'dp_template_set_origin': (i32, [vp, cp]),
        'dp_template_origin': (cp, [vp]),
        'dp_template_free': (None, [vp]),
        'dp_template_get_info': (i32, [vp, P(dp_template_info)]),
        'dp_template_record': (i32, [vp, u32, P(u64), P(u64), P(cp), P(cp)]),
        'dp_template_unpack': (i32, [vp, vp, u32, u64, u64, i32, C.c_void_p]),
        'dp_scan': (i32, [vp, vp, u32, P(cp), P(C.c_int32), u64, u64, i32, i32,
                          P(vp), P(u64), P(dp_scan_stats)]),
        'dp_tables_load': (i32, [vp, P(dp_nn_tables)]),
        'dp_duplex_dg': (i32, [vp, vp, C.c_void_p, u64, u32, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_void_p, P(dp_conditions), P(vp), P(u64), P(dp_dg_stats)]),
        'dp_find': (i32, [vp, vp, u32, P(cp), P(C.c_int32), u64, u64, i32, i32, C.c_void_p, C.c_void_p,
                          C.c_void_p, P(dp_conditions), P(vp), P(u64), P(vp), P(u64),
                          P(dp_scan_stats), P(dp_dg_stats)]),
        'dp_find_begin': (i32, [vp, vp, u32, P(cp), P(C.c_int32), u64, u64, i32, i32, C.c_void_p, C.c_void_p,
                                C.c_void_p, P(dp_conditions)]),
        'dp_find_end': (i32, [vp, P(vp), P(u64), P(vp), P(u64), P(dp_scan_stats), P(dp_dg_stats)]),
        'dp_set_wire': (i32, [vp, C.c_void_p, u64]),
        'dp_set_timing': (i32, [vp, i32]),
        'dp_find_again': (i32, [vp, vp, P(vp), P(u64), vp]),
        'dp_scan_prepare_kernels': (i32, [vp, u32, P(C.c_char_p), P(i32), i32]),
        'dp_jit_selftest': (i32, [u32, P(C.c_char_p), P(i32), i32, C.c_char_p, C.c_char_p]),
        'dp_jit_describe': (i32, [u32, P(C.c_char_p), P(i32), i32, C.c_char_p, u64]),
        'dp_comm_export': (i32, [vp, i32, i32, u32, C.c_char_p]),
        'dp_comm_attach': (i32, [vp, C.c_char_p]),
        'dp_comm_destroy': (i32, [vp]),
        'dp_comm_gathered': (i32, [vp, i32, P(vp), P(u64), P(i32)]),
        'dp_comm_hits': (i32, [vp, i32, P(vp), P(u64)]),
        'dp_comm_set_layout': (i32, [vp, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
        'dp_stream': (vp, [vp]),
        'dp_synchronize': (i32, [vp]),
        'dp_hit_windows': (i32, [vp, vp, C.c_void_p, u64, u32, C.c_void_p, C.c_void_p]),
        'dp_calibrate': (i32, [vp, i32, i32, P(C.c_double), P(C.c_double)]),
        'dp_variant_tm': (i32, [vp, u32, u32, C.c_void_p, C.c_double, C.c_double, P(dp_conditions), C.c_void_p]),
        'dp_variant_tm_static': (i32, [vp, u32, u32, C.c_void_p, C.c_double, C.c_double, P(dp_conditions), C.c_void_p]),
        'dp_variant_tm_static_limits': (i32, [P(u32), P(u64), P(u64)]),
        'dp_pair_products': (i32, [vp, u32, C.c_void_p, u32, C.c_void_p, C.c_int64, C.c_int64, P(vp), P(u64), C.c_void_p]),
        'dp_dimer_dg': (i32, [vp, u32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_void_p, C.c_void_p, P(dp_conditions), C.c_void_p]),
        'dp_hairpin_tables_load': (i32, [vp, P(dp_hairpin_tables)]),
        'dp_find_hairpins': (i32, [vp, u32, C.c_void_p, C.c_void_p, C.c_void_p, u32, C.c_void_p, C.c_void_p,
                                   C.c_void_p, P(dp_conditions), C.c_double, C.c_void_p, P(C.c_float)]),
        'dp_find_dimers': (i32, [vp, u32, C.c_void_p, C.c_void_p, C.c_void_p, u32, C.c_void_p, C.c_void_p,
                                 P(dp_conditions), u32, C.c_void_p, P(C.c_float)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return sig


EXPORTS = ('dp_abi_version dp_device_count dp_create dp_create_error dp_destroy dp_last_error dp_free '
           'dp_abort_requested dp_device_info dp_template_from_records dp_template_from_ascii '
           'dp_template_from_fasta dp_template_set_record_name dp_template_save dp_template_load dp_template_image_size dp_template_save_mem dp_template_load_mem dp_template_set_origin dp_template_origin dp_template_free dp_template_get_info dp_template_record '
           'dp_template_unpack dp_scan dp_tables_load dp_duplex_dg dp_find dp_find_begin dp_find_end '
           'dp_set_wire dp_set_timing dp_find_again dp_scan_prepare_kernels dp_jit_selftest dp_jit_describe dp_comm_export dp_comm_attach dp_comm_destroy dp_comm_gathered dp_comm_hits dp_comm_set_layout dp_stream dp_synchronize '
           'dp_hit_windows dp_variant_tm dp_variant_tm_static dp_variant_tm_static_limits dp_pair_products dp_dimer_dg dp_find_dimers dp_hairpin_tables_load dp_find_hairpins dp_calibrate').split()


def load():
    """dlopen the library (no CUDA call is made).  Raises if it is missing."""
    global _lib
    with _lock:
        if _lib is None:
            from .build import build_native, _stale
            if _LIB_OVERRIDE:
                lib = C.CDLL(_LIB_OVERRIDE)
                _declare(lib)
                _lib = lib
                return _lib
            if _stale():              # sources newer than the .so: rebuild (seconds) or fail loudly
                build_native()
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    'libdegenprimer_b200.so is not built (%s). Run `python -m degenprimer_b200.build`; '
                    'there is no CPU fallback.' % LIB_PATH)
            lib = C.CDLL(LIB_PATH)
            _declare(lib)
            if lib.dp_abi_version() != 1:
                raise RuntimeError('libdegenprimer_b200.so has ABI %d, expected 1' % lib.dp_abi_version())
            _lib = lib
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def jit_plan(masters, mismatches, strands=DP_STRAND_BOTH):
    """dp_jit_describe: the prefilter plan of the specialised scan kernel of a primer set —
    [dict(pattern, K, need, p_alive, groups=[[window positions], ...])], mode.  No GPU needed."""
    lib = load()
    n = len(masters)
    m = (C.c_char_p * n)(*[s.encode('ascii') for s in masters])
    if isinstance(mismatches, int):
        mismatches = [mismatches] * n
    k = (C.c_int32 * n)(*[int(x) for x in mismatches])
    buf = C.create_string_buffer(1 << 20)
    rc = lib.dp_jit_describe(n, m, k, int(strands), buf, len(buf))
    if rc < 0:
        raise NativeError(rc, lib.dp_last_error(None).decode(errors='replace'))
    plans, mode = [], None
    for line in buf.value.decode().splitlines():
        f = line.split()
        if f[1] == 'mode':
            mode = f[2]
            continue
        plans.append(dict(pattern=int(f[2]), K=int(f[4]), need=int(f[6]), p_alive=float(f[8]),
                          groups=[[int(x) for x in g.split(',')] for g in f[10:]]))
    return plans, mode


class Template(object):
    """A set of records resident in HBM (handle owned by this object)."""

    def __init__(self, ctx, handle, keepalive=None):
        self.ctx = ctx
        self._h = handle
        self._keep = keepalive
        info = dp_template_info()
        ctx._check(ctx.lib.dp_template_get_info(handle, C.byref(info)))
        self.n_records = info.n_records
        self.total_bases = info.total_bases
        self.padded_bases = info.padded_bases
        self.n_exception_runs = info.n_exception_runs
        self.device_bytes = info.device_bytes
        self.starts, self.lengths, self.ids, self.descriptions = [], [], [], []
        s, n, i, d = C.c_uint64(), C.c_uint64(), C.c_char_p(), C.c_char_p()
        for r in range(self.n_records):
            ctx._check(ctx.lib.dp_template_record(handle, r, C.byref(s), C.byref(n), C.byref(i), C.byref(d)))
            self.starts.append(s.value)
            self.lengths.append(n.value)
            self.ids.append(i.value.decode())
            self.descriptions.append(d.value.decode())

    def unpack(self, record=0, begin=0, n=None, reverse=False):
        if n is None:
            n = self.lengths[record] - begin
        buf = C.create_string_buffer(max(1, n))
        self.ctx._check(self.ctx.lib.dp_template_unpack(self.ctx._h, self._h, record, begin, n,
                                                        1 if reverse else 0, buf))
        return buf.raw[:n].decode('ascii')

    def set_record_names(self, ids, descriptions):
        """Name the records (dp_template_set_record_name): for loaders of formats other than FASTA."""
        for r, (rid, descr) in enumerate(zip(ids, descriptions)):
            self.ctx._check(self.ctx.lib.dp_template_set_record_name(self._h, r, rid.encode(), descr.encode()))
        self.ids, self.descriptions = list(ids), list(descriptions)

    def save(self, path):
        """dp_template_save: planes + run table + record table as they sit in HBM."""
        self.ctx._check(self.ctx.lib.dp_template_save(self.ctx._h, self._h, os.fsencode(path)))
        return path

    @property
    def image_bytes(self):
        n = C.c_uint64()
        self.ctx._check(self.ctx.lib.dp_template_image_size(self._h, C.byref(n)))
        return n.value

    def save_image(self, address, capacity):
        """dp_template_save_mem: the packed-file image into caller-owned host memory
        (pin it, and Context.template_load_image brings it back with three DMA transfers)."""
        self.ctx._check(self.ctx.lib.dp_template_save_mem(self.ctx._h, self._h, C.c_void_p(address), int(capacity)))
        return self.image_bytes

    @property
    def origin(self):
        return (self.ctx.lib.dp_template_origin(self._h) or b'').decode(errors='replace')

    @origin.setter
    def origin(self, text):
        self.ctx._check(self.ctx.lib.dp_template_set_origin(self._h, text.encode()))

    def close(self):
        if self._h is not None and self.ctx._h is not None:
            self.ctx.lib.dp_template_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _Owner(object):
    """Releases a library-allocated result block when the numpy view over it dies."""
    __slots__ = ('lib', 'ptr')

    def __init__(self, lib, ptr):
        self.lib, self.ptr = lib, ptr

    def __del__(self):
        try:
            self.lib.dp_free(C.c_void_p(self.ptr))
        except Exception:
            pass


class Query(object):
    """Masters, mismatch limits and unambiguous variants of a set of primers, laid out
    for dp_find / dp_duplex_dg (variant v of primer p = len(p) bytes at
    offset[p] + v*len(p) of one blob)."""

    def __init__(self, masters, mismatches, variants):
        n = len(masters)
        self.n = n
        self.masters = [s if isinstance(s, str) else s.decode('ascii') for s in masters]
        self.m = (C.c_char_p * n)(*[s.encode('ascii') for s in self.masters])
        if isinstance(mismatches, int):
            mismatches = [mismatches] * n
        self.k = (C.c_int32 * n)(*[int(x) for x in mismatches])
        self.lens = np.array([len(s) for s in self.masters], dtype=np.uint8)
        self.vc = np.array([len(v) for v in variants], dtype=np.uint32)
        sizes = self.vc.astype(np.uint64) * self.lens.astype(np.uint64)
        self.vo = np.concatenate(([0], np.cumsum(sizes)[:-1])).astype(np.uint64)
        self.vb = ''.join(''.join(v) for v in variants).encode('ascii')
        self.vc_p, self.vo_p = _ptr(self.vc), _ptr(self.vo)
        self.vb_p = C.cast(C.c_char_p(self.vb), C.c_void_p)
        self.n_variants = int(self.vc.sum())


class Context(object):
    """One CUDA context + stream on one device."""

    def __init__(self, device=0):
        self.lib = load()
        h = self.lib.dp_create(int(device))
        if not h:
            raise NativeError(DP_ERR_CUDA, self.lib.dp_create_error().decode())
        self._h = h
        self.device = int(device)
        self._tables_key = None

    def close(self):
        if self._h is not None:
            self.lib.dp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc == DP_OK:
            return
        msg = self.lib.dp_last_error(self._h).decode(errors='replace')
        if rc == DP_ERR_ABORTED:
            raise Aborted(msg)
        if rc == DP_ERR_LENGTH:
            raise ValueError('SearchEngine._find: template sequence should be longer or equal to primer '
                             'sequence and both should be grater than zero. (%s)' % msg)
        if rc == DP_ERR_ALPHABET:
            raise KeyError(msg)
        raise NativeError(rc, msg)

    def abort(self, flag=True):
        return self.lib.dp_abort_requested(self._h, 1 if flag else 0)

    def device_info(self):
        name = C.create_string_buffer(256)
        sm, maj, mnr, khz = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        mem = C.c_uint64()
        self._check(self.lib.dp_device_info(self._h, name, C.byref(sm), C.byref(maj), C.byref(mnr),
                                            C.byref(khz), C.byref(mem)))
        return dict(name=name.value.decode(), sm_count=sm.value, cc=(maj.value, mnr.value),
                    clock_khz=khz.value, hbm_bytes=mem.value)

    # ---- templates -------------------------------------------------------------
    def template_from_sequences(self, seqs):
        """seqs: list of bytes / str / (address, length) pairs."""
        keep, ptrs, lens = [], [], []
        for s in seqs:
            if isinstance(s, tuple):
                ptrs.append(int(s[0]))
                lens.append(int(s[1]))
                continue
            if isinstance(s, str):
                s = s.encode('latin-1')
            elif not isinstance(s, bytes):
                s = bytes(s)
            keep.append(s)
            ptrs.append(C.cast(C.c_char_p(s), C.c_void_p).value or 0)
            lens.append(len(s))
        n = len(ptrs)
        arr = (C.c_void_p * max(1, n))(*ptrs)
        ln = (C.c_uint64 * max(1, n))(*lens)
        out = C.c_void_p()
        self._check(self.lib.dp_template_from_records(self._h, n, arr, ln, C.byref(out)))
        return Template(self, out.value)

    def template_from_fasta(self, path):
        out = C.c_void_p()
        self._check(self.lib.dp_template_from_fasta(self._h, os.fsencode(path), C.byref(out)))
        return Template(self, out.value)

    def template_load(self, path):
        """dp_template_load: a packed template written by Template.save."""
        out = C.c_void_p()
        self._check(self.lib.dp_template_load(self._h, os.fsencode(path), C.byref(out)))
        return Template(self, out.value)

    def template_load_image(self, address, nbytes, verify=True):
        """dp_template_load_mem: a packed image (Template.save_image, or a cache file read or
        mmap()ed into memory) -> resident template.  verify=False skips the checksum pass."""
        out = C.c_void_p()
        self._check(self.lib.dp_template_load_mem(self._h, C.c_void_p(address), int(nbytes), 1 if verify else 0,
                                                  C.byref(out)))
        return Template(self, out.value)

    # ---- scan ---------------------------------------------------------------------
    def scan(self, template, masters, mismatches, seg_begin=0, seg_end=2 ** 64 - 1,
             strands=DP_STRAND_BOTH, algo=DP_ALGO_AUTO):
        """Returns (hits structured array, stats dict)."""
        n = len(masters)
        m = (C.c_char_p * n)(*[s.encode('ascii') if isinstance(s, str) else s for s in masters])
        if isinstance(mismatches, int):
            mismatches = [mismatches] * n
        k = (C.c_int32 * n)(*[int(x) for x in mismatches])
        hits, n_hits, st = C.c_void_p(), C.c_uint64(), dp_scan_stats()
        self._check(self.lib.dp_scan(self._h, template._h, n, m, k, seg_begin, seg_end, strands, algo,
                                     C.byref(hits), C.byref(n_hits), C.byref(st)))
        arr = self._take(hits, n_hits.value, HIT_DTYPE)
        return arr, st

    def _take(self, ptr, n, dtype):
        """A malloc()ed result array as numpy.  Small results are copied and released;
        large ones are wrapped in place and released (dp_free) when the last view dies."""
        nbytes = n * dtype.itemsize
        if nbytes < (1 << 16):
            try:
                if n:      # string_at + bytearray: two small copies, a third of the cost of frombuffer().copy()
                    return np.frombuffer(bytearray(C.string_at(ptr.value, nbytes)), dtype=dtype)
                return np.zeros(0, dtype=dtype)
            finally:
                self.lib.dp_free(ptr)
        buf = (C.c_char * nbytes).from_address(ptr.value)
        buf._owner = _Owner(self.lib, ptr.value)          # numpy -> memoryview -> buf -> owner
        return np.frombuffer(buf, dtype=dtype)

    def prepare(self, masters, mismatches, variants, jit=None, strands=DP_STRAND_BOTH):
        """The primer side of dp_find as ready-made C arrays (build once, scan many
        templates).  variants: one list of unambiguous sequences per master.
        jit: also compile the scan kernel specialised for these primers
        (dp_scan_prepare_kernels, ~0.3 s once per process and primer set).  It pays for
        a query that is scanned thousands of times or over Gbp templates (a screening
        loop); default off, DEGENPRIMER_B200_JIT=1 turns it on for every prepare().
        Without NVRTC the generic kernel serves."""
        q = Query(masters, mismatches, variants)
        if jit is None:
            jit = os.environ.get('DEGENPRIMER_B200_JIT', '0') == '1'
        q.jit = False
        if jit:
            rc = self.lib.dp_scan_prepare_kernels(self._h, q.n, q.m, q.k, int(strands))
            if rc == 0:
                q.jit = True
            elif rc != DP_ERR_STATE:
                self._check(rc)
            else:
                q.jit_error = self.lib.dp_last_error(self._h).decode(errors='replace')
        return q

    def find(self, template, masters, mismatches=None, variants=None, cond=None, seg_begin=0,
             seg_end=2 ** 64 - 1, strands=DP_STRAND_BOTH, algo=DP_ALGO_AUTO):
        """Fused scan + duplex dG (dp_find).  `masters` is a Query from prepare(), or
        the masters themselves together with `mismatches` and `variants`.
        Returns (hits, dimers, scan stats, dG stats)."""
        q = masters if isinstance(masters, Query) else Query(masters, mismatches, variants)
        hits, n_hits, dimers, n_dimers = C.c_void_p(), C.c_uint64(), C.c_void_p(), C.c_uint64()
        st, dst = dp_scan_stats(), dp_dg_stats()
        self._check(self.lib.dp_find(self._h, template._h, q.n, q.m, q.k, seg_begin, seg_end, strands, algo,
                                     q.vc_p, q.vo_p, q.vb_p, C.byref(cond),
                                     C.byref(hits), C.byref(n_hits), C.byref(dimers), C.byref(n_dimers),
                                     C.byref(st), C.byref(dst)))
        h = self._take(hits, n_hits.value, HIT_DTYPE)
        d = self._take(dimers, n_dimers.value, DIMER_DTYPE)
        return h, d, st, dst

    def find_begin(self, template, query, cond, seg_begin=0, seg_end=2 ** 64 - 1,
                   strands=DP_STRAND_BOTH, algo=DP_ALGO_AUTO):
        """First half of find(): every kernel is queued on `stream`, nothing is waited for."""
        q = query
        self._check(self.lib.dp_find_begin(self._h, template._h, q.n, q.m, q.k, seg_begin, seg_end, strands,
                                           algo, q.vc_p, q.vo_p, q.vb_p, C.byref(cond)))

    def find_end(self):
        """Second half of find(): waits for the stream and returns (hits, dimers, stats, dG stats)."""
        hits, n_hits, dimers, n_dimers = C.c_void_p(), C.c_uint64(), C.c_void_p(), C.c_uint64()
        st, dst = dp_scan_stats(), dp_dg_stats()
        self._check(self.lib.dp_find_end(self._h, C.byref(hits), C.byref(n_hits), C.byref(dimers),
                                         C.byref(n_dimers), C.byref(st), C.byref(dst)))
        h = self._take(hits, n_hits.value, HIT_DTYPE)
        d = self._take(dimers, n_dimers.value, DIMER_DTYPE)
        return h, d, st, dst

    def find_again(self, cond):
        """The duplex stage of the last find() of this context under other conditions: the hits
        stay on the device, `dimers['hit']` indexes the hits that find() returned.
        Returns (dimers, dG stats)."""
        dimers, n_dimers = C.c_void_p(), C.c_uint64()
        dst = dp_dg_stats()
        self._check(self.lib.dp_find_again(self._h, C.byref(cond), C.byref(dimers), C.byref(n_dimers), C.byref(dst)))
        return self._take(dimers, n_dimers.value, DIMER_DTYPE), dst

    def set_timing(self, on):
        """Per-kernel CUDA-event timing of every call (ms_* stats); off saves ~15 us of host time."""
        self._check(self.lib.dp_set_timing(self._h, 1 if on else 0))

    def set_wire(self, device_ptr, nbytes):
        """Device buffer that receives {header, raw hits} of every call (multi-GPU gather)."""
        self._check(self.lib.dp_set_wire(self._h, C.c_void_p(device_ptr or 0), int(nbytes)))

    # ---- multi-GPU exchange (peer-memory stores from the publish kernel) -----------
    def comm_export(self, world, rank, max_hits=4096):
        """Allocates this rank's receive buffers; returns their 128 bytes of IPC handles."""
        buf = C.create_string_buffer(128)
        self._check(self.lib.dp_comm_export(self._h, int(world), int(rank), int(max_hits), buf))
        self._gather_world, self._gather_max = int(world), int(max_hits)
        return buf.raw

    def comm_attach(self, all_handles):
        """`all_handles`: the 128-byte blobs of comm_export of every rank, in rank order."""
        blob = b''.join(all_handles)
        if len(blob) != 128 * self._gather_world:
            raise ValueError('expected %d handle blobs of 128 bytes' % self._gather_world)
        self._check(self.lib.dp_comm_attach(self._h, blob))

    def comm_destroy(self):
        self._check(self.lib.dp_comm_destroy(self._h))

    def comm_set_layout(self, base, local_start, offset, record, record_len):
        """dp_comm_set_layout: where every rank's local records were cut from, so that
        gathered_hits() returns whole-genome coordinates.  base=None switches it off."""
        if base is None:
            self._check(self.lib.dp_comm_set_layout(self._h, None, None, None, None, None))
            return
        b = np.ascontiguousarray(base, dtype=np.uint32)
        ls = np.ascontiguousarray(local_start, dtype=np.uint64)
        of = np.ascontiguousarray(offset, dtype=np.uint64)
        rc = np.ascontiguousarray(record, dtype=np.uint32)
        rl = np.ascontiguousarray(record_len, dtype=np.uint64)
        if len(b) != self._gather_world + 1 or not (len(ls) == len(of) == len(rc) == len(rl) == int(b[-1])):
            raise ValueError('layout tables do not match the world size / entry count')
        self._check(self.lib.dp_comm_set_layout(self._h, _ptr(b), _ptr(ls), _ptr(of), _ptr(rc), _ptr(rl)))

    def gathered_hits(self, age=0):
        """Hits of every rank for a call (one segment per rank: record = sender rank,
        positions in that segment), sorted.  age 0 = the newest call (waits for the
        slowest rank), 1 = the call before (waits for this rank's stream), 2 = two calls
        back (already on the host: collect it between find_begin and find_end).
        None if a slot overflowed or a rank holds a multi-record template."""
        hits, n = C.c_void_p(), C.c_uint64()
        rc = self.lib.dp_comm_hits(self._h, int(age), C.byref(hits), C.byref(n))
        if rc == DP_ERR_UNSUPPORTED:
            return None
        self._check(rc)
        return self._take(hits, n.value, HIT_DTYPE)

    def gathered(self, age=0):
        """(raw wire records of every rank for a call, in one array; sender rank of
        every record), or None when some rank had more hits than a slot holds."""
        ptr, slot_b, w = C.c_void_p(), C.c_uint64(), C.c_int()
        self._check(self.lib.dp_comm_gathered(self._h, int(age), C.byref(ptr), C.byref(slot_b), C.byref(w)))
        world, slot, cap = w.value, slot_b.value, self._gather_max
        raw = (C.c_uint8 * (slot * world)).from_address(ptr.value)
        buf = np.frombuffer(raw, dtype=np.uint8).reshape(world, slot)
        head = buf[:, :64].copy().view(np.uint64)
        counts = (head[:, 0] & np.uint64(0xFFFFFFFF)).astype(np.int64)
        if int(counts.max()) > cap or (head[:, 4] & np.uint64(2)).any():     # a slot overflowed or is truncated
            return None
        recs = buf[:, WIRE_HEADER_BYTES:].view(WIRE_DTYPE)            # (world, cap) view of pinned memory
        keep = np.arange(cap)[None, :] < counts[:, None]
        return recs[keep], np.repeat(np.arange(world), counts)

    @property
    def stream(self):
        """cudaStream_t of this context as an integer (for torch.cuda.ExternalStream)."""
        return int(self.lib.dp_stream(self._h) or 0)

    def synchronize(self):
        self._check(self.lib.dp_synchronize(self._h))

    # ---- thermodynamics ----------------------------------------------------------
    def load_tables(self, tables, key=None):
        if key is not None and key == self._tables_key:
            return
        self._check(self.lib.dp_tables_load(self._h, C.byref(tables)))
        self._tables_key = key

    def duplex_dg(self, template, hits, primer_lens, var_counts, var_offsets, variants, cond):
        """hits: HIT_DTYPE array grouped by primer.  Returns (DIMER_DTYPE array, stats)."""
        hits = np.ascontiguousarray(hits, dtype=HIT_DTYPE)
        pl = np.ascontiguousarray(primer_lens, dtype=np.uint8)
        vc = np.ascontiguousarray(var_counts, dtype=np.uint32)
        vo = np.ascontiguousarray(var_offsets, dtype=np.uint64)
        vb = variants if isinstance(variants, bytes) else variants.encode('ascii')
        out, n_out, st = C.c_void_p(), C.c_uint64(), dp_dg_stats()
        self._check(self.lib.dp_duplex_dg(self._h, template._h, _ptr(hits), len(hits), len(pl), _ptr(pl),
                                          _ptr(vc), _ptr(vo), C.cast(C.c_char_p(vb), C.c_void_p),
                                          C.byref(cond), C.byref(out), C.byref(n_out), C.byref(st)))
        arr = self._take(out, n_out.value, DIMER_DTYPE)
        return arr, st

    def pair_products(self, fwd_pos, rev_pos, min_amplicon, max_amplicon):
        """dp_pair_products: the pairing loop of MixtureFactory.create_PCR_mixture on the device.
        Returns (pairs as an (n, 2) uint32 array of (forward index, reverse index) in the reference's
        order, rev_start per forward site)."""
        f = np.ascontiguousarray(fwd_pos, dtype=np.int64)
        r = np.ascontiguousarray(rev_pos, dtype=np.int64)
        rs = np.zeros(len(f), dtype=np.uint32)
        out, n = C.c_void_p(), C.c_uint64()
        self._check(self.lib.dp_pair_products(self._h, len(f), _ptr(f), len(r), _ptr(r), int(min_amplicon),
                                              int(max_amplicon), C.byref(out), C.byref(n), _ptr(rs)))
        pairs = self._take(out, n.value, PAIR_DTYPE)
        return pairs, rs

    def calibrate(self, which, reps=5):
        """Lane-ops/s of a pipe: 'lop3', 'popc' or 'lds'."""
        idx = {'lop3': 0, 'popc': 1, 'lds': 2}[which]
        rate, ms = C.c_double(), C.c_double()
        self._check(self.lib.dp_calibrate(self._h, idx, reps, C.byref(rate), C.byref(ms)))
        return rate.value

    def hit_windows(self, template, hits, primer_lens):
        """List of window strings, one per hit (strand orientation)."""
        hits = np.ascontiguousarray(hits, dtype=HIT_DTYPE)
        pl = np.ascontiguousarray(primer_lens, dtype=np.uint8)
        out = np.zeros((len(hits), 64), dtype=np.uint8)
        if len(hits):
            self._check(self.lib.dp_hit_windows(self._h, template._h, _ptr(hits), len(hits), len(pl),
                                                _ptr(pl), _ptr(out)))
        lens = pl[hits['primer']] if len(hits) else []
        raw = out.tobytes()
        return [raw[64 * i:64 * i + int(n)].decode('ascii') for i, n in enumerate(lens)]

    def variant_tm(self, seqs, concentration, cond, conversion_degree=0.5):
        n = len(seqs)
        if n == 0:
            return np.zeros(0)
        L = len(seqs[0])
        blob = ''.join(seqs).encode('ascii')
        tm = np.zeros(n, dtype=np.float64)
        self._check(self.lib.dp_variant_tm(self._h, n, L, C.cast(C.c_char_p(blob), C.c_void_p),
                                           float(concentration), float(conversion_degree),
                                           C.byref(cond), _ptr(tm)))
        return tm

    def dimer_dg(self, seqs1, seqs2, masks, offsets, cond):
        n = len(seqs1)
        s1 = ''.join(seqs1).encode('ascii')
        s2 = ''.join(seqs2).encode('ascii')
        l1 = np.array([len(s) for s in seqs1], dtype=np.uint8)
        l2 = np.array([len(s) for s in seqs2], dtype=np.uint8)
        o1 = np.concatenate(([0], np.cumsum(l1, dtype=np.uint64)[:-1])).astype(np.uint64)
        o2 = np.concatenate(([0], np.cumsum(l2, dtype=np.uint64)[:-1])).astype(np.uint64)
        mk = np.ascontiguousarray(masks, dtype=np.uint64)
        of = np.ascontiguousarray(offsets, dtype=np.int32)
        out = np.zeros(n, dtype=np.float64)
        self._check(self.lib.dp_dimer_dg(self._h, n, C.cast(C.c_char_p(s1), C.c_void_p), _ptr(o1), _ptr(l1),
                                         C.cast(C.c_char_p(s2), C.c_void_p), _ptr(o2), _ptr(l2),
                                         _ptr(mk), _ptr(of), C.byref(cond), _ptr(out)))
        return out


    def load_hairpin_tables(self, tables, key=None):
        if key is not None and key == getattr(self, '_hp_tables_key', None):
            return
        self._check(self.lib.dp_hairpin_tables_load(self._h, C.byref(tables)))
        self._hp_tables_key = key

    def find_hairpins(self, seqs, seed_seq, seed_mask, seed_offset, cond, max_hairpin_dG):
        """dp_find_hairpins: returns (BEST_STRUCTURE_DTYPE[n_seeds], kernel ms)."""
        ln = np.array([len(s) for s in seqs], dtype=np.uint8)
        if any(len(s) > 64 for s in seqs):
            raise ValueError('sequences longer than 64 nt are not supported by dp_find_hairpins')
        off = np.concatenate(([0], np.cumsum(ln, dtype=np.uint64)[:-1])).astype(np.uint64)
        blob = ''.join(seqs).encode('ascii')
        sq = np.ascontiguousarray(seed_seq, dtype=np.uint32)
        mk = np.ascontiguousarray(seed_mask, dtype=np.uint64)
        so = np.ascontiguousarray(seed_offset, dtype=np.int32)
        out = np.zeros(len(sq), dtype=BEST_STRUCTURE_DTYPE)
        ms = C.c_float(0.0)
        self._check(self.lib.dp_find_hairpins(self._h, len(seqs), C.cast(C.c_char_p(blob), C.c_void_p), _ptr(off),
                                              _ptr(ln), len(sq), _ptr(sq), _ptr(mk), _ptr(so), C.byref(cond),
                                              float(max_hairpin_dG), _ptr(out), C.byref(ms)))
        return out, float(ms.value)

    def find_dimers(self, seqs, pairs, cond):
        """dp_find_dimers: `seqs` list of str, `pairs` (n, 2) indices into it.  Returns
        (records[n_pairs, stride] of FRONT_DIMER_DTYPE, kernel ms)."""
        pairs = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
        n_pairs = len(pairs)
        ln = np.array([len(s) for s in seqs], dtype=np.uint8)
        if any(len(s) > 64 for s in seqs):
            raise ValueError('sequences longer than 64 nt are not supported by dp_find_dimers')
        off = np.concatenate(([0], np.cumsum(ln, dtype=np.uint64)[:-1])).astype(np.uint64)
        blob = ''.join(seqs).encode('ascii')
        a = np.ascontiguousarray(pairs[:, 0])
        b = np.ascontiguousarray(pairs[:, 1])
        stride = int((ln[a].astype(np.int64) + ln[b].astype(np.int64) - 1).max()) if n_pairs else 1
        out = np.zeros((n_pairs, stride), dtype=FRONT_DIMER_DTYPE)
        ms = C.c_float(0.0)
        self._check(self.lib.dp_find_dimers(self._h, len(seqs), C.cast(C.c_char_p(blob), C.c_void_p), _ptr(off),
                                            _ptr(ln), n_pairs, _ptr(a), _ptr(b), C.byref(cond), stride,
                                            _ptr(out), C.byref(ms)))
        return out, float(ms.value)


PACK_MAGIC = b'DPB200PK'
PACK_HEADER = np.dtype([('magic', 'S8'), ('version', '<u4'), ('n_records', '<u4'), ('n_words', '<u8'),
                        ('n_words_alloc', '<u8'), ('total_bases', '<u8'), ('n_runs', '<u8'),
                        ('names_bytes', '<u8'), ('checksum', '<u8')])
PACK_RECORD = np.dtype([('start', '<u8'), ('len', '<u8'), ('id_len', '<u4'), ('descr_len', '<u4')])
PACK_RUN = np.dtype([('start', '<u8'), ('end', '<u8'), ('letter', '<u4'), ('pad', '<u4')])
assert PACK_HEADER.itemsize == 64 and PACK_RECORD.itemsize == 24 and PACK_RUN.itemsize == 24


def pack_checksum(*chunks):
    """The checksum of a packed-template file (dp_pack_sum in csrc/dp_template.cu): the header with its
    checksum field zeroed and the six sections (record table, names + origin, run table, plane lo, plane hi,
    plane exc), each zero-padded to a multiple of 8 bytes, form one sequence of little-endian 64-bit words; word i goes to lane i & 3 (h_l starts at
    SEED + l; h = (h ^ w) * P; h ^= h >> 29) and the four lanes are folded at the end."""
    mask = (1 << 64) - 1
    prime = 0x100000001B3
    hs = [(0xCBF29CE484222325 + lane) & mask for lane in range(4)]
    i = 0
    for chunk in chunks:
        b = bytes(chunk)
        for k in range(0, len(b), 8):
            w = int.from_bytes(b[k:k + 8].ljust(8, b'\0'), 'little')
            h = ((hs[i & 3] ^ w) * prime) & mask
            hs[i & 3] = h ^ (h >> 29)
            i += 1
    r = hs[0]
    for lane in range(1, 4):
        r = ((r * prime) & mask) ^ hs[lane]
    return r


def packed_info(path):
    """Header and record table of a packed-template file, read on the host (no GPU): a dict with
    n_records, total_bases, n_runs, records = [(id, description, start, len)], expected_bytes, bytes."""
    size = os.path.getsize(path)
    with open(path, 'rb') as fh:
        raw = fh.read(64)
        if len(raw) < 64:
            raise ValueError('%s: shorter than a packed-template header' % path)
        h = np.frombuffer(raw, dtype=PACK_HEADER)[0]
        if bytes(h['magic']) != PACK_MAGIC:
            raise ValueError('%s: not a packed template (bad magic)' % path)
        if int(h['version']) != 3:
            raise ValueError('%s: packed-template version %d, expected 3' % (path, int(h['version'])))
        n = int(h['n_records'])
        recs = np.frombuffer(fh.read(24 * n), dtype=PACK_RECORD)
        names = fh.read(int(h['names_bytes']))
    want = 64 + 24 * n + int(h['names_bytes']) + 24 * int(h['n_runs']) + 12 * int(h['n_words_alloc'])
    out, p = [], 0
    for r in recs:
        rid = names[p:p + int(r['id_len'])].decode()
        p += int(r['id_len'])
        descr = names[p:p + int(r['descr_len'])].decode()
        p += int(r['descr_len'])
        out.append((rid, descr, int(r['start']), int(r['len'])))
    origin = names[p:].decode(errors='replace')
    return {'origin': origin, 'n_records': n, 'total_bases': int(h['total_bases']), 'n_words': int(h['n_words']),
            'n_runs': int(h['n_runs']), 'checksum': int(h['checksum']), 'records': out,
            'expected_bytes': want, 'bytes': size}


_contexts = {}


def default_device():
    return int(os.environ.get('DEGENPRIMER_B200_DEVICE', os.environ.get('LOCAL_RANK', '0')))


def context(device=None):
    """Process-wide context per device, created on first use (after any fork)."""
    if device is None:
        device = default_device()
    key = (os.getpid(), device)
    ctx = _contexts.get(key)
    if ctx is None:
        ctx = Context(device)
        _contexts[key] = ctx
    return ctx
