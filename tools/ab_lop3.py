#!/usr/bin/env python3
"""Explicit-LOP3 prefilter against the plain-C one on an eighth of config 4 (one GPU, a few seconds):
scan kernel time of both, and the hit set of each against the generic bit-slice kernel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from degenprimer_b200 import TD_Functions as tdf          # noqa: E402
from degenprimer_b200 import synthetic                    # noqa: E402
from degenprimer_b200.SeqUtils import unambiguous_sequences   # noqa: E402

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.125
wl = synthetic.make_workload('config4_20plex_3Gbp', scale=scale)
primers, k = wl['primers'], wl['k']
var = [unambiguous_sequences(p) for p in primers]
tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0))
ctx = tdf._ctx()
cond = tdf.conditions()
tmpl = ctx.template_from_sequences([(r.ctypes.data, len(r)) for r in wl['records']])
ref, _ = ctx.scan(tmpl, primers, k, algo=2)
print('generic kernel: %d hits' % len(ref), flush=True)
for mode in ('0', '1', '0', '1'):
    os.environ['DEGENPRIMER_B200_JIT_LOP3'] = mode
    c = tdf.new_context()                  # prepared queries are cached per context: a fresh one per mode
    q = c.prepare(primers, k, var, jit=True)
    ms = []
    for i in range(6):
        hits, dimers, st, dst = c.find(tmpl, q, cond=cond)
        ms.append(st.ms_scan)
    print('lop3=%s scan %.4f ms (best of 6, algo %d) hits %d equal-to-generic %s dimers %d' % (
        mode, min(ms[1:]), st.algo, len(hits), bool(np.array_equal(hits, ref)), len(dimers)), flush=True)
