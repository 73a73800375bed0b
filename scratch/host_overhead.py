import sys, time, os; sys.path.insert(0, '.')
import numpy as np, torch
from degenprimer_b200 import TD_Functions as tdf
from degenprimer_b200.synthetic import make_workload
from degenprimer_b200.SeqUtils import unambiguous_sequences
wl = make_workload('config2_tp53_chr17', scale=float(os.environ.get('SCALE', '1.0')))
tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0))
ctx = tdf._ctx()
t = ctx.template_from_sequences([wl['records'][0].tobytes()])
var=[unambiguous_sequences(p) for p in wl['primers']]
q = ctx.prepare(wl['primers'], 1, var)
cond = tdf.conditions()
for _ in range(5): ctx.find(t, q, cond=cond)
N=int(os.environ.get('N','20'))
t0=time.perf_counter()
for _ in range(N): r = ctx.find(t, q, cond=cond)
dt=(time.perf_counter()-t0)/N*1e6
print('find wall %.1f us; device ev0->ev5 %.1f us; scan %.1f main %.1f dg %.1f' % (dt, r[2].ms_total*1e3, r[2].ms_scan*1e3, r[2].ms_main*1e3, r[3].ms_kernel*1e3))
