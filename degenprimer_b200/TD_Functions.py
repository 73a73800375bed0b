"""Thermodynamic entry points with the reference's names, computed on the GPU.

Host mirror of DegenPrimer/TD_Functions.py: module globals `NN`, `PCR_P`,
`AcquireParameters()` (:45-58) and the functions the rest of DegenPrimer calls
(`primer_template_Tr/Tm` :60-101, `dimer_dG_corrected` :124-206,
`equilibrium_constant` :279-283, `primer_DNA_conversion_degree` :286-297).
Every dG / Tm is evaluated by kernels in csrc/dp_dg.cu through the C-ABI; the
two closed-form scalar helpers at the bottom are one `exp` / one `sqrt`.
"""
import threading
from contextlib import contextmanager
from copy import deepcopy
from math import exp, nextafter, sqrt, inf, log

from . import _native as nat
from .PCR_Parameters import PCR_Parameters
from .UnifiedNN import UnifiedNN

NN = UnifiedNN()
PCR_P = PCR_Parameters()
_PCR_P = None

max_dimer_dG = -3      # DegenPrimer/SecStructures.py:30
min_K = 100            # DegenPrimer/SecStructures.py:32


_acquire_lock = threading.RLock()


@contextmanager
def AcquireParameters():
    """Work on a private snapshot of PCR_P (TD_Functions.py:50-58).  The reference runs one
    process per worker; here worker threads share the module, so other threads wait while
    one holds the snapshot (nesting in one thread still asserts, as in the reference)."""
    global _PCR_P, PCR_P
    with _acquire_lock:
        assert _PCR_P is None, 'TD_Functions: PCR parameters are already acquired.'
        try:
            _PCR_P, PCR_P = PCR_P, deepcopy(PCR_P)
            yield
        finally:
            PCR_P, _PCR_P = _PCR_P, None


def equilibrium_constant(dG_T, T):
    """TD_Functions.py:279-283."""
    return exp(-1000 * dG_T / (NN.R * (T - NN.K0)))


def max_dG_for_min_K(T, K_min=None):
    """Largest double dG with equilibrium_constant(dG, T) >= K_min.

    Duplex._recalculate drops a dimer when `K < min_K` (SecStructures.py:303-304).
    K is monotone in dG, so the kernel applies that filter as `dG <= cutoff`;
    the cutoff is located here with the very same `exp` expression, which makes
    the device-side filter agree with the reference on every double.
    """
    K_min = min_K if K_min is None else K_min
    g = -NN.R * (T - NN.K0) * log(K_min) / 1000.0
    for _ in range(64):                       # walk to the exact boundary
        if equilibrium_constant(g, T) >= K_min:
            up = nextafter(g, inf)
            if equilibrium_constant(up, T) >= K_min:
                g = up
                continue
            return g
        g = nextafter(g, -inf)
    raise RuntimeError('max_dG_for_min_K did not converge')


def conditions(pcr_p=None):
    """dp_conditions block for the current PCR_P."""
    p = PCR_P if pcr_p is None else pcr_p
    c = nat.dp_conditions()
    c.PCR_T = p.PCR_T
    c.Na_eq_log = p.Na_eq_log
    c.max_dimer_dG = max_dimer_dG
    c.max_dG_for_K = max_dG_for_min_K(p.PCR_T)
    c.DNA = p.DNA
    c.DMSO = p.DMSO
    return c


def _ctx():
    """Process-wide native context with the current NN tables on the device."""
    ctx = nat.context()
    if ctx._tables_key != UnifiedNN._version:
        ctx.load_tables(UnifiedNN.device_tables(), key=UnifiedNN._version)
    return ctx


def _ctx_with_hairpin_tables():
    ctx = _ctx()
    ctx.load_hairpin_tables(UnifiedNN.hairpin_device_tables(), key=UnifiedNN._version)
    return ctx


def new_context(device=None):
    """A further native context on a device (own stream, workspaces and result buffers) with
    the current NN tables loaded: one per call that is to be in flight on the GPU at the same
    time, or one per worker thread.  Templates may be shared between the contexts of a device."""
    ctx = nat.Context(nat.default_device() if device is None else device)
    ctx.load_tables(UnifiedNN.device_tables(), key=UnifiedNN._version)
    return ctx


def primer_template_Tr(sequence, concentration, conversion_degree):
    """TD_Functions.py:60-98, two-state model, on the device."""
    if not (0 < conversion_degree < 1):
        raise ValueError('TD_Functions.primer_template_Tr: equilibrium ratio should be in the (0;1) interval.')
    return float(_ctx().variant_tm([str(sequence)], concentration, conditions(), conversion_degree)[0])


def primer_template_Tm(sequence, concentration):
    return primer_template_Tr(sequence, concentration, 0.5)


def primer_template_Tms(sequences, concentration):
    """All variants of a primer in one launch (Primer.calculate_Tms, Primer.py:121-140)."""
    return _ctx().variant_tm([str(s) for s in sequences], concentration, conditions())


def dimer_dG_corrected(dimer, seq1, seq2):
    """TD_Functions.py:124-206 for one dimer, incl. dangling ends (offset != 0)."""
    return float(dimers_dG_corrected([dimer], [seq1], [seq2])[0])


def dimers_dG_corrected(dimers, seqs1, seqs2):
    """Batch form: one thread per dimer."""
    masks = []
    for d in dimers:
        m = 0
        for i in d.fwd_matches:
            m |= 1 << i
        masks.append(m)
    out = _ctx().dimer_dg([str(s) for s in seqs1], [str(s) for s in seqs2], masks,
                          [d.offset for d in dimers], conditions())
    if (out != out).any():
        raise ValueError('UnifiedNN: a nearest-neighbour key of this dimer is not in the database')
    return out


def primer_DNA_conversion_degree(primer_concentration, K):
    """TD_Functions.py:286-297 (closed form, used by reports)."""
    P = primer_concentration
    D = PCR_P.DNA
    _b = (K * P + K * D + 1)
    disc = _b * _b - 4 * K * (K * P * D)
    DUP = (_b - sqrt(disc)) / (2 * K)
    return DUP / min(P, D)
