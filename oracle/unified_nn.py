"""ORACLE (test infrastructure, see oracle/README.md) — Unified NN tables.

Restates DegenPrimer/UnifiedNN.py:126-139 (constants), :208-221 (table load
incl. reversed-key aliases), :224-246 (stack lookup with ambiguous-letter
averaging + caching), :248-252 (dangling ends), :254-276 (internal loops),
:317-319 (loop dS).  Data: oracle/tables.json (made by make_tables.py from the
reference's data/*.csv).
"""
import json
import os
from math import log

from .sequtils import unambiguous_sequences

R = 1.9872                    # UnifiedNN.py:127
K0 = -273.15                  # :128
K37 = 37 - K0                 # :129
DS_NA = 0.368                 # :132
T_DMSO = 0.75                 # :133
LOOP_COEFF = 2.44             # :134
IK37 = -1000.0 / K37          # :136
LRK37 = LOOP_COEFF * R * K37 / 1000.0   # :137
TERMINAL_MISMATCH = (-1.23 + -0.21) / 2  # :140


class Pars(object):
    """dH/dS/dG triple (UnifiedNN.py:41-88)."""
    __slots__ = ('dH', 'dS', 'dG')

    def __init__(self, dH=0, dS=0, dG=0):
        self.dH, self.dS, self.dG = dH, dS, dG

    def __repr__(self):
        return 'Pars(dH=%r, dS=%r, dG=%r)' % (self.dH, self.dS, self.dG)


def _load():
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, 'tables.json')) as fh:
        doc = json.load(fh)
    stacks = {k: Pars(*v) for k, v in doc['stacks'].items()}
    # UnifiedNN.py:214-215 — every key also answers to its reversed string
    for k in list(stacks):
        stacks[k[::-1]] = stacks[k]
    dangling = {k: Pars(*v) for k, v in doc['dangling'].items()}
    loops = {int(k): {'internal': v[0], 'hairpin': v[1]} for k, v in doc['loops'].items()}
    tri_tetra = {k: {'dH': v[0], 'dG': v[1]} for k, v in doc['tri_tetra_hairpin_loops'].items()}
    return stacks, dangling, loops, tri_tetra


STACKS, DANGLING, LOOPS, TRI_TETRA = _load()
INI = STACKS['ini']
TER = STACKS['ter']
SYM = STACKS['sym']


def int_dPar_37(seq, rev):
    """UnifiedNN.py:224-246.  `seq` 5'->3' over `rev` 3'->5', two letters each."""
    key = seq + '/' + rev
    hit = STACKS.get(key)
    if hit is not None:
        return hit
    seqs = unambiguous_sequences(seq)
    revs = unambiguous_sequences(rev)
    if len(seqs) == 1 and len(revs) == 1:
        raise ValueError('stack not in the table: %s' % key)
    acc = Pars()
    for s in seqs:
        for r in revs:
            p = int_dPar_37(s, r)
            acc.dH += p.dH
            acc.dS += p.dS
            acc.dG += p.dG
    n = float(len(seqs) * len(revs))
    acc.dH /= n
    acc.dS /= n
    acc.dG /= n
    STACKS[key] = acc            # cached under both spellings, as the reference does
    STACKS[key[::-1]] = acc
    return acc


def end_dPar_37(seq, rev):
    """UnifiedNN.py:248-252."""
    return DANGLING[seq + '/' + rev]


def _extrapolate_loop_dG_37(length, kind):
    """UnifiedNN.py:254-261."""
    if length < 30:
        base = length
        while base not in LOOPS:
            base += 1
    else:
        base = 30
    return LOOPS[base][kind] + LRK37 * log(float(length) / base)


def internal_loop_dG_37(length):
    """UnifiedNN.py:264-276 (the >30 warning print is dropped)."""
    if length in LOOPS:
        g = LOOPS[length]['internal']
    else:
        g = _extrapolate_loop_dG_37(length, 'internal')
    g += 2 * TERMINAL_MISMATCH
    return g


def internal_loop_dS_37(seq):
    """UnifiedNN.py:317-319 — depends on len(seq) only."""
    return internal_loop_dG_37(len(seq) - 2) * IK37


def hairpin_loop_dG_37(seq):
    """UnifiedNN.py:279-307; seq includes the closing pair (the >30 warning print is dropped)."""
    length = len(seq) - 2
    if length in LOOPS:
        g = LOOPS[length]['hairpin']
        if length == 3:
            if seq in TRI_TETRA:
                g += TRI_TETRA[seq]['dG']
            if seq[0] in 'AT':
                g += 0.5
        elif length == 4:
            if seq in TRI_TETRA:
                g += TRI_TETRA[seq]['dG']
            g += TERMINAL_MISMATCH
        else:
            g += TERMINAL_MISMATCH
        return g
    g = _extrapolate_loop_dG_37(length, 'hairpin')
    g += TERMINAL_MISMATCH
    return g


def hairpin_loop_dH_37(seq):
    """UnifiedNN.py:310-315."""
    if len(seq) - 2 > 4:
        return 0
    try:
        return TRI_TETRA[seq]['dH']
    except KeyError:
        return 0


def hairpin_loop_dS_37(seq):
    """UnifiedNN.py:321-325."""
    return (hairpin_loop_dG_37(seq) - hairpin_loop_dH_37(seq)) * IK37
