"""ORACLE (test infrastructure, see oracle/README.md) — PCR conditions.

Restates DegenPrimer/PCR_Parameters.py:66-97 (defaults, Na_eq, Na_eq_log) and
the module-global `PCR_P` + `AcquireParameters` snapshot of
DegenPrimer/TD_Functions.py:45-58.  Concentrations are mol/L, DMSO is %, PCR_T
is Celsius.
"""
from contextlib import contextmanager
from copy import copy
from math import log, sqrt

FIELDS = ('Mg', 'Na', 'dNTP', 'DNA', 'DMSO', 'PCR_T', 'Na_eq', 'Na_eq_log')


class Conditions(object):
    def __init__(self, **kw):
        self.Mg = 1.5e-3       # PCR_Parameters.py:72-77
        self.Na = 50e-3
        self.dNTP = 0.1e-3
        self.DNA = 5e-9
        self.DMSO = 0.0
        self.PCR_T = 60.0
        self.set(kw)

    def set(self, params):
        """PCR_Parameters.py:52-56: unknown names are ignored, then recalc."""
        for name in params:
            if name in FIELDS:
                setattr(self, name, params[name])
        self._recalculate()

    def _recalculate(self):
        """PCR_Parameters.py:81-87 (von Ahsen 2001 divalent-cation correction)."""
        self.Na_eq = self.Na + sqrt((self.Mg - self.dNTP) * 14.4)
        self.Na_eq_log = log(self.Na_eq)


PCR_P = Conditions()
_saved = None


@contextmanager
def AcquireParameters():
    """TD_Functions.py:50-58: work on a private snapshot of PCR_P."""
    global PCR_P, _saved
    assert _saved is None, 'parameters are already acquired'
    try:
        _saved, PCR_P = PCR_P, copy(PCR_P)
        yield
    finally:
        PCR_P, _saved = _saved, None
