"""ORACLE (test infrastructure, see oracle/README.md) — duplex thermodynamics.

Restates DegenPrimer/TD_Functions.py:
  primer_template_Tr / primer_template_Tm   :60-101
  dimer_dG_corrected                        :124-206
  equilibrium_constant                      :279-283
The order of every floating-point addition is kept, because the downstream
filter (dG > -3, K < 100; SecStructures.py:302-304) is a knife edge.
"""
from math import exp, log

from . import conditions as cond
from . import unified_nn as nn
from .sequtils import COMPLEMENT


def primer_template_Tr(sequence, concentration, conversion_degree):
    """TD_Functions.py:60-98.  `sequence` is a str (5'->3')."""
    if not (0 < conversion_degree < 1):
        raise ValueError('equilibrium ratio should be in the (0;1) interval.')
    seq = str(sequence)
    # Bio.Seq.complement(): not reversed, so rev[n] pairs with seq[n]
    rev = ''.join(COMPLEMENT[c] for c in seq)
    P = concentration
    D = cond.PCR_P.DNA
    DUP = conversion_degree * min(P, D)
    K = DUP / ((P - DUP) * (D - DUP))
    dH = nn.INI.dH
    dS = nn.INI.dS
    if seq[0] in 'AT':
        dH += nn.TER.dH
        dS += nn.TER.dS
    if seq[-1] in 'AT':
        dH += nn.TER.dH
        dS += nn.TER.dS
    for n in range(len(seq) - 1):
        p = nn.int_dPar_37(seq[n:n + 2], rev[n:n + 2])
        dH += p.dH
        dS += p.dS
    dS = dS + nn.DS_NA * len(seq) * cond.PCR_P.Na_eq_log
    return nn.K0 + dH * 1000 / (dS - nn.R * log(K)) - nn.T_DMSO * cond.PCR_P.DMSO


def primer_template_Tm(sequence, concentration):
    """TD_Functions.py:100-101."""
    return primer_template_Tr(sequence, concentration, 0.5)


def dimer_dG_corrected(dimer, seq1, seq2):
    """TD_Functions.py:124-206.  seq1, seq2 both 5'->3'."""
    fm = dimer.fwd_matches
    off = dimer.offset
    s = str(seq1)
    s_len = len(s)
    r = str(seq2)[::-1]          # now 3'->5', index-aligned with s via `off`
    r_len = len(r)
    dG = 0
    dH = nn.INI.dH
    dS = nn.INI.dS
    # ---- left end (:140-158)
    f0 = fm[0]
    if f0 == 0 and off < 0:                       # 3' dangling
        r0 = f0 - off
        p = nn.end_dPar_37(r[r0] + r[r0 - 1], s[0] + '-')
        dH += p.dH
        dS += p.dS
    elif f0 == 0 and off == 0:                    # blunt: terminal AT penalty
        if s[0] in 'AT':
            dH += nn.TER.dH
            dS += nn.TER.dS
    elif f0 == off:                               # 5' dangling
        p = nn.end_dPar_37(s[f0 - 1] + s[f0], '-' + r[0])
        dH += p.dH
        dS += p.dS
    elif f0 > off:                                # terminal mismatch
        dG += nn.TERMINAL_MISMATCH
    # ---- right end (:160-178)
    f1 = fm[-1]
    r1 = f1 - off
    if f1 == s_len - 1 and r1 < r_len - 1:        # 5' dangling
        p = nn.end_dPar_37(r[r1 + 1] + r[r1], '-' + s[-1])
        dH += p.dH
        dS += p.dS
    elif r1 == r_len - 1 and f1 < s_len - 1:      # 3' dangling
        p = nn.end_dPar_37(s[f1] + s[f1 + 1], r[-1] + '-')
        dH += p.dH
        dS += p.dS
    elif f1 < s_len - 1 and r1 < r_len - 1:       # terminal mismatch
        dG += nn.TERMINAL_MISMATCH
    elif f1 == s_len - 1 and r1 == r_len - 1:     # blunt
        if s[-1] in 'AT':
            dH += nn.TER.dH
            dS += nn.TER.dS
    # ---- stacks, single/double mismatches and internal loops (:180-202)
    for i in range(dimer.num_matches - 1):
        a = fm[i]
        b = fm[i + 1]
        ra = a - off
        rb = b - off
        gap = b - a
        if gap < 4:
            p = nn.int_dPar_37(s[a:a + 2], r[ra:ra + 2])
            dH += p.dH
            dS += p.dS
            if gap > 1:
                p = nn.int_dPar_37(r[rb - 1:rb + 1][::-1], s[b - 1:b + 1][::-1])
                dH += p.dH
                dS += p.dS
        else:
            dS += nn.internal_loop_dS_37(s[a:b + 1])
    # ---- salt (:204) and temperature (:205)
    dS += nn.DS_NA * (f1 - f0) * cond.PCR_P.Na_eq_log
    return dG + dH - (cond.PCR_P.PCR_T - nn.K0) * dS / 1000.0


def equilibrium_constant(dG_T, T):
    """TD_Functions.py:279-283."""
    return exp(-1000 * dG_T / (nn.R * (T - nn.K0)))


def hairpin_dG_corrected(hairpin, seq):
    """TD_Functions.py:209-276.  hairpin.fwd_matches ascending, hairpin.rev_matches descending,
    fwd_matches[i] pairs with rev_matches[i]."""
    fm, rm = hairpin.fwd_matches, hairpin.rev_matches
    s = str(seq)
    s_len = len(s)
    dG = 0
    dH = nn.INI.dH
    dS = nn.INI.dS
    f0, r0 = fm[0], rm[0]
    if f0 == 0 and r0 < s_len - 1:                 # 3' dangling (:228-232)
        p = nn.end_dPar_37(s[r0] + s[r0 + 1], s[0] + '-')
        dH += p.dH
        dS += p.dS
    elif f0 > 0 and r0 == s_len - 1:               # 5' dangling (:233-237)
        p = nn.end_dPar_37(s[f0 - 1] + s[f0], '-' + s[-1])
        dH += p.dH
        dS += p.dS
    elif f0 > 0 and r0 < s_len - 1:                # terminal mismatch
        dG += nn.TERMINAL_MISMATCH
    elif f0 == 0 and r0 == s_len - 1:              # blunt
        if s[0] in 'AT':
            dH += nn.TER.dH
            dS += nn.TER.dS
    for i in range(len(fm) - 1):                   # :246-267
        a, b = fm[i], fm[i + 1]
        ra, rb = rm[i], rm[i + 1]
        gap = b - a
        if gap < 4:
            p = nn.int_dPar_37(s[a:a + 2], s[ra - 1:ra + 1][::-1])
            dH += p.dH
            dS += p.dS
            if gap > 1:
                p = nn.int_dPar_37(s[rb:rb + 2], s[b - 1:b + 1][::-1])
                dH += p.dH
                dS += p.dS
        else:
            dS += nn.internal_loop_dS_37(s[a:b + 1])
    f1 = fm[-1]
    hp = s[f1:rm[-1] + 1]                          # :269-272
    dH += nn.hairpin_loop_dH_37(hp)
    dS += nn.hairpin_loop_dS_37(hp)
    dS += nn.DS_NA * (f1 - f0) * cond.PCR_P.Na_eq_log
    return dG + dH - (cond.PCR_P.PCR_T - nn.K0) * dS / 1000.0
