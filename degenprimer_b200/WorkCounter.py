"""Hierarchical weighted progress counter (duck type of DegenPrimer/WorkCounter.py:41-181).

The hot path only reports coarse progress (a kernel launch is one unit of work),
but the object keeps the reference's interface: `set_work`, `set_subwork`,
`add_subwork`, `count`, `done`, `__getitem__`, `total`, `progress`, `ratio`,
`percent`, `changed_percent`.
"""
from time import time


class WorkCounter(object):
    def __init__(self, work=1, weights=None, parent=None, subwork=False):
        self._parent = parent
        self._subcounters = []
        self._last_ratio = None
        self._changed = False
        self.set_work(work, weights, subwork)

    def __bool__(self):
        return True

    def __len__(self):
        return len(self._subcounters)

    def __getitem__(self, index):
        return self._subcounters[index]

    def set_parent(self, parent):
        self._parent = parent

    def set_work(self, work, weights=None, subwork=False):
        self.reset()
        self._subcounters = []
        self._weights = []
        if hasattr(work, '__len__'):
            work = len(work)
        assert work > 0, 'Counter: %x: work size should be grater than zero' % id(self)
        if subwork:
            self._work = 0
            self._own_work = 0
            self._subwork = 0
            self.add_subwork(work, weights)
        else:
            self._work = work
            self._own_work = work
            self._subwork = 0

    def set_subwork(self, work, weights=None):
        self.set_work(work, weights=weights, subwork=True)

    def add_subwork(self, work=1, weights=None):
        weights = list(weights) if weights else [1] * work
        assert len(weights) == work, 'a weight should be provided for every subwork.'
        assert min(weights) >= 0, 'Weights should be >= 0.'
        total = sum(weights)
        self._work += total
        self._subwork += total
        self._weights.extend(weights)
        for _ in range(work):
            self._subcounters.append(WorkCounter(parent=self))

    def elapsed(self):
        return time() - self._time0

    def last_checked(self):
        return time() - self._time_1

    def changed(self):
        return self._changed

    def count(self, num=1):
        assert self._count + num <= self._own_work, 'the work counted exceeds total work amount'
        self._count += num

    def done(self):
        self._count = self._work - self._subwork

    def reset(self):
        for c in self._subcounters:
            c.reset()
        self._count = 0
        self._time0 = time()
        self._time_1 = self._time0

    def total(self):
        return self._work

    def progress(self):
        return sum(c.ratio() * w for w, c in zip(self._weights, self._subcounters)) + self._count

    def ratio(self):
        ratio = self.progress() / float(self._work) if self._work else 1.0
        self._changed = ratio != self._last_ratio
        self._last_ratio = ratio
        return ratio

    def percent(self):
        self._time_1 = time()
        return self.ratio() * 100.0

    def changed_percent(self):
        percent = self.percent()
        return percent if self._changed else None
