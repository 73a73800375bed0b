import sys, time; sys.path.insert(0, '.')
import numpy as np, torch
from degenprimer_b200 import TD_Functions as tdf
from degenprimer_b200.synthetic import make_workload
from degenprimer_b200.SeqView import ResidentRecord
from degenprimer_b200.SearchEngine import SearchEngine
from degenprimer_b200.Primer import Primer
from degenprimer_b200.SeqUtils import Seq, SeqRecord, unambiguous_sequences
from degenprimer_b200.WorkCounter import WorkCounter
wl = make_workload('config2_tp53_chr17')
tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0))
ctx = tdf._ctx()
host = torch.from_numpy(wl['records'][0]).pin_memory()
dev = torch.empty_like(host, device='cuda')
for _ in range(3): dev.copy_(host, non_blocking=True); torch.cuda.synchronize()
t0=time.perf_counter()
for _ in range(5): dev.copy_(host, non_blocking=True); torch.cuda.synchronize()
dt=(time.perf_counter()-t0)/5
print('torch pinned H2D: %.3f ms  %.1f GB/s' % (dt*1e3, host.numel()/dt/1e9))
srcs=[(host.data_ptr(), host.numel())]
def T(f, n=10):
    f(); f(); f()
    t0=time.perf_counter()
    for _ in range(n): r=f()
    return (time.perf_counter()-t0)/n*1e3, r
ms, t = T(lambda: ctx.template_from_sequences(srcs))
print('template_from_sequences (pinned): %.3f ms' % ms)
ms2, _ = T(lambda: ctx.template_from_sequences(srcs).close())
print('  + close: %.3f ms' % ms2)
gp=[]
for i,p in enumerate(wl['primers']):
    g=Primer(SeqRecord(Seq(p), id='p%d'%i), 1e-6); g.generate_components(); gp.append(g)
se=SearchEngine(); rec=ResidentRecord(t,0)
ms, m = T(lambda: se.find_matches(WorkCounter(), rec, gp[0], 1))
print('find_matches: %.3f ms' % ms, se.last_stats['scan'])
ms, r = T(lambda: se.compile_duplexes_mp(WorkCounter(), *m))
print('compile_duplexes_mp: %.3f ms' % ms, se.last_stats['dg'])
ms, r = T(lambda: se.find_duplexes(WorkCounter(), rec, gp[0], 1))
print('find_duplexes (fused): %.3f ms' % ms, se.last_stats)
ms, r = T(lambda: ctx.scan(t, wl['primers'], 1))
print('ctx.scan both primers: %.3f ms' % ms, r[1])
hits=r[0]
var=[unambiguous_sequences(p) for p in wl['primers']]
ms, r = T(lambda: ctx.duplex_dg(t, hits, [20,20],[1,1],[0,20], ''.join(wl['primers']), tdf.conditions()))
print('ctx.duplex_dg: %.3f ms' % ms, r[1])
ms, r = T(lambda: ctx.find(t, wl['primers'], 1, var, tdf.conditions()))
print('ctx.find fused: %.3f ms' % ms, r[2], r[3])
