"""PCR conditions — host mirror of DegenPrimer/PCR_Parameters.py:66-97.

Same attribute names, defaults and `set()` behaviour.  The reference keeps the
values in multiprocessing.Value cells so forked workers see updates; here the
GPU is the worker, so plain floats suffice and `deepcopy` gives the private
snapshot `TD_Functions.AcquireParameters` needs.
"""
from math import log, sqrt


class PCR_Parameters(object):
    parameters = ['Mg', 'Na', 'dNTP', 'DNA', 'DMSO', 'PCR_T', 'Na_eq', 'Na_eq_log']

    def __init__(self):
        self.Mg = 1.5e-3       # M
        self.Na = 50e-3        # M
        self.dNTP = 0.1e-3     # M
        self.DNA = 5e-9        # M, template concentration
        self.DMSO = 0.0        # %, v/v
        self.PCR_T = 60.0      # C
        self._recalculate()

    def set(self, params):
        for name in params:
            if name in self.parameters:
                setattr(self, name, params[name])
        self._recalculate()

    def _recalculate(self):
        """Divalent cation correction (von Ahsen et al. 2001), PCR_Parameters.py:81-87.
        Mg < dNTP raises ValueError from sqrt, as in the reference."""
        self.Na_eq = self.Na + sqrt((self.Mg - self.dNTP) * 14.4)
        self.Na_eq_log = log(self.Na_eq)

    def __deepcopy__(self, memo):
        c = PCR_Parameters.__new__(PCR_Parameters)
        c.__dict__.update(self.__dict__)
        return c

    def __str__(self):
        w = max(len(p) for p in self.parameters)
        return '\n' + ''.join('%s%s: %f\n' % (n, ' ' * (w - len(n)), getattr(self, n))
                              for n in self.parameters)
