"""Unified nearest-neighbour tables, flattened for the device.

Host-side mirror of DegenPrimer/UnifiedNN.py: the same class attributes and
lookups (`internal_NN`, `dangling_ends`, `loops`, `ini/ter/sym`, `int_dPar_37`,
`end_dPar_37`, `internal_loop_dG_37`, `internal_loop_dS_37`, constants
:126-139), used only to BUILD the `dp_nn_tables` block the kernels index
(include/degenprimer_b200.h).  No dG is computed here.
"""
import csv
import os
from math import log

from . import _native as nat
from . import nn_data
from .SeqUtils import unambiguous_sequences


class Pars(object):
    """dH / dS / dG triple (DegenPrimer/UnifiedNN.py:41-88)."""
    __slots__ = ['dH', 'dS', 'dG']

    def __init__(self, dH=0, dS=0, dG=0):
        self.dH, self.dS, self.dG = dH, dS, dG

    def __str__(self):
        return ', '.join('%s=%f' % (s, getattr(self, s)) for s in self.__slots__)

    __repr__ = __str__


class Loops(object):
    __slots__ = ['internal', 'hairpin']

    def __init__(self, internal=0, hairpin=0):
        self.internal, self.hairpin = internal, hairpin


_COMP = {'A': 'T', 'T': 'A', 'G': 'C', 'C': 'G'}


class UnifiedNN(object):
    R = 1.9872
    K0 = -273.15
    K37 = 37 - K0
    dG_Na_coefficient_oligo = -0.114
    dG_Na_coefficient_poly = -0.175
    dS_Na_coefficient = 0.368
    T_DMSO_coefficient = 0.75
    Loop_coefficient = 2.44
    iK37 = -1000.0 / K37
    LRK37 = Loop_coefficient * R * K37 / 1000.0
    Terminal_mismatch_mean = (-1.23 + -0.21) / 2

    internal_NN = None
    dangling_ends = None
    loops = None
    tri_tetra_hairpin_loops = None
    ter = Pars()
    ini = Pars()
    sym = Pars()
    _inited = False
    _version = 0          # bumped on every (re)load: device copies are refreshed

    def __init__(self):
        if not UnifiedNN._inited:
            self.load_tables()

    def __bool__(self):
        return self._inited

    @staticmethod
    def _read_csv(path):
        """The reference's TAB-separated, '"'-quoted layout (UnifiedNN.py:163-199)."""
        with open(path, newline='') as fh:
            rows = list(csv.reader(fh, delimiter='\t', quotechar='"'))
        head = rows[0][1:]
        return {r[0]: dict(zip(head, map(float, r[1:]))) for r in rows[1:] if r}

    @classmethod
    def load_tables(cls, directory=None):
        """Built-in tables, or the reference's csv files found in `directory`."""
        if directory is None:
            stacks = {k: Pars(*v) for k, v in nn_data.STACKS.items()}
            ends = {k: Pars(*v) for k, v in nn_data.DANGLING_ENDS.items()}
            loops = {k: Loops(*v) for k, v in nn_data.LOOPS.items()}
            tri_tetra = {k: Pars(dH=v[0], dG=v[1]) for k, v in nn_data.TRI_TETRA_HAIRPIN_LOOPS.items()}
        else:
            stacks = {k: Pars(**v) for k, v in cls._read_csv(os.path.join(directory, 'internal-NN.csv')).items()}
            ends = {k: Pars(**v) for k, v in cls._read_csv(os.path.join(directory, 'dangling-ends.csv')).items()}
            loops = {int(k): Loops(**v) for k, v in cls._read_csv(os.path.join(directory, 'loops.csv')).items()}
            tri_tetra = {k: Pars(**v) for k, v in
                         cls._read_csv(os.path.join(directory, '3-4-hairpin-loops.csv')).items()}
        for k in list(stacks):            # UnifiedNN.py:214-215
            stacks[k[::-1]] = stacks[k]
        cls.internal_NN, cls.dangling_ends, cls.loops = stacks, ends, loops
        cls.tri_tetra_hairpin_loops = tri_tetra
        cls.ter = stacks.get('ter', Pars())
        cls.ini = stacks.get('ini', Pars())
        cls.sym = stacks.get('sym', Pars())
        cls._inited = True
        cls._version += 1

    @classmethod
    def int_dPar_37(cls, seq, rev):
        """UnifiedNN.py:224-246 (table lookup; expansion average for ambiguous letters)."""
        key = seq + '/' + rev
        pars = cls.internal_NN.get(key)
        if pars is not None:
            return pars
        seqs = unambiguous_sequences(seq)
        revs = unambiguous_sequences(rev)
        if len(seqs) == 1 and len(revs) == 1:
            raise ValueError('UnifiedNN.int_dPar_37: sequence is not in the database: %s' % key)
        pars = Pars()
        for s in seqs:
            for r in revs:
                p = cls.int_dPar_37(s, r)
                pars.dH += p.dH
                pars.dS += p.dS
                pars.dG += p.dG
        n = float(len(seqs) * len(revs))
        pars.dH /= n
        pars.dS /= n
        pars.dG /= n
        cls.internal_NN[key] = pars
        cls.internal_NN[key[::-1]] = pars
        return pars

    @classmethod
    def end_dPar_37(cls, seq, rev):
        return cls.dangling_ends[seq + '/' + rev]

    @classmethod
    def _extrapolate_loop_dG_37(cls, length, loop_type):
        if length < 30:
            exp_len = length
            while exp_len not in cls.loops:
                exp_len += 1
        else:
            exp_len = 30
        return getattr(cls.loops[exp_len], loop_type) + cls.LRK37 * log(float(length) / exp_len)

    @classmethod
    def internal_loop_dG_37(cls, length):
        try:
            loop_dG = cls.loops[length].internal
        except KeyError:
            loop_dG = cls._extrapolate_loop_dG_37(length, 'internal')
        return loop_dG + 2 * cls.Terminal_mismatch_mean

    @classmethod
    def internal_loop_dS_37(cls, seq):
        return cls.internal_loop_dG_37(len(seq) - 2) * cls.iK37

    @classmethod
    def dG_T(cls, dH, dS, temperature):
        return dH - (temperature - cls.K0) * dS / 1000.0

    @classmethod
    def hairpin_loop_dG_37(cls, seq):
        """UnifiedNN.py:279-307; seq includes the closing pair."""
        length = len(seq) - 2
        if length in cls.loops:
            loop_dG = cls.loops[length].hairpin
            if length == 3:
                if seq in cls.tri_tetra_hairpin_loops:
                    loop_dG += cls.tri_tetra_hairpin_loops[seq].dG
                if seq[0] in 'AT':
                    loop_dG += 0.5
            elif length == 4:
                if seq in cls.tri_tetra_hairpin_loops:
                    loop_dG += cls.tri_tetra_hairpin_loops[seq].dG
                loop_dG += cls.Terminal_mismatch_mean
            else:
                loop_dG += cls.Terminal_mismatch_mean
            return loop_dG
        return cls._extrapolate_loop_dG_37(length, 'hairpin') + cls.Terminal_mismatch_mean

    @classmethod
    def hairpin_loop_dH_37(cls, seq):
        """UnifiedNN.py:310-315."""
        if len(seq) - 2 > 4:
            return 0
        try:
            return cls.tri_tetra_hairpin_loops[seq].dH
        except KeyError:
            return 0

    @classmethod
    def hairpin_loop_dS_37(cls, seq):
        return (cls.hairpin_loop_dG_37(seq) - cls.hairpin_loop_dH_37(seq)) * cls.iK37

    # ---- device block --------------------------------------------------------------
    @classmethod
    def hairpin_device_tables(cls):
        """dp_hairpin_tables filled exactly as dp_hairpin_dg_device indexes it."""
        if not cls._inited:
            cls.load_tables()
        t = nat.dp_hairpin_tables()
        nan = float('nan')
        for n in range(64):
            if n < 3:
                t.loop_dG[n] = nan
            elif n in cls.loops:
                t.loop_dG[n] = cls.loops[n].hairpin
            else:
                t.loop_dG[n] = cls._extrapolate_loop_dG_37(n, 'hairpin')
        for i in range(1024):
            t.tri_dH[i] = t.tri_dG[i] = nan
        for i in range(4096):
            t.tetra_dH[i] = t.tetra_dG[i] = nan
        for key, p in cls.tri_tetra_hairpin_loops.items():
            idx = 0
            for c in key:
                idx = idx * 4 + nat.BASES.index(c)
            if len(key) == 5:
                t.tri_dH[idx], t.tri_dG[idx] = p.dH, p.dG
            elif len(key) == 6:
                t.tetra_dH[idx], t.tetra_dG[idx] = p.dH, p.dG
            else:
                raise ValueError('3-4-hairpin-loops: key %r is neither a tri- nor a tetra-loop' % key)
        t.iK37 = cls.iK37
        return t

    @classmethod
    def device_tables(cls):
        """dp_nn_tables filled exactly as the kernels index it."""
        if not cls._inited:
            cls.load_tables()
        t = nat.dp_nn_tables()
        nan = float('nan')
        B, Lt = nat.BASES, nat.LETTERS
        for xi, x in enumerate(B):
            for yi, y in enumerate(B):
                for wi, w in enumerate(Lt):
                    i = (xi * 4 + yi) * 15 + wi
                    p = cls.int_dPar_37(x + y, _COMP[x] + w)             # TD_Functions.py:187-191
                    t.stack_dH[i], t.stack_dS[i] = p.dH, p.dS
                    p = cls.int_dPar_37(_COMP[y] + w, y + x)             # TD_Functions.py:193-197
                    t.rstack_dH[i], t.rstack_dS[i] = p.dH, p.dS
        for kind, fmt in enumerate(('%s%s/-%s', '%s%s/%s-')):
            for xi, x in enumerate(B):
                for yi, y in enumerate(B):
                    for zi, z in enumerate(B):
                        i = ((kind * 4 + xi) * 4 + yi) * 4 + zi
                        p = cls.dangling_ends.get(fmt % (x, y, z))
                        t.dangling_dH[i] = p.dH if p else nan
                        t.dangling_dS[i] = p.dS if p else nan
        for n in range(64):
            t.loop_dS[n] = cls.internal_loop_dG_37(n) * cls.iK37 if n >= 1 else nan
        t.ini_dH, t.ini_dS = cls.ini.dH, cls.ini.dS
        t.ter_dH, t.ter_dS = cls.ter.dH, cls.ter.dS
        t.terminal_mismatch = cls.Terminal_mismatch_mean
        t.dS_Na = cls.dS_Na_coefficient
        t.R = cls.R
        t.K0 = cls.K0
        t.T_DMSO = cls.T_DMSO_coefficient
        return t
