import os, sys, time; sys.path.insert(0, '.')
import numpy as np, torch, torch.distributed as dist
rank=int(os.environ['RANK']); world=int(os.environ['WORLD_SIZE']); local=int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
os.environ['DEGENPRIMER_B200_DEVICE']=str(local)
from degenprimer_b200 import TD_Functions as tdf
from degenprimer_b200.synthetic import make_workload
from degenprimer_b200.SeqUtils import unambiguous_sequences
from degenprimer_b200.distributed import HitExchange
wl = make_workload('config2_tp53_chr17', seed_offset=1000*rank)
tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0))
ctx = tdf._ctx(); cond = tdf.conditions()
t = ctx.template_from_sequences([wl['records'][0].tobytes()])
var=[unambiguous_sequences(p) for p in wl['primers']]
q = ctx.prepare(wl['primers'], 1, var)
ex = HitExchange(ctx, max_hits=1024)
AGE=int(os.environ.get('AGE','1'))
T={k:0.0 for k in 'begin end collect'.split()}
N=200
dev=[]
for it in range(N+10):
    if it==10:
        for k in T: T[k]=0.0
        dist.barrier(); torch.cuda.synchronize(); t00=time.perf_counter()
    a=time.perf_counter(); ctx.find_begin(t,q,cond)
    b=time.perf_counter(); r=ctx.find_end()
    c=time.perf_counter(); h=ex.collect_hits(AGE) if it>0 else None
    d=time.perf_counter()
    for k,v in zip(T,(b-a,c-b,d-c)): T[k]+=v
    dev.append((r[2].ms_scan, r[3].ms_kernel, r[2].ms_total))
tot=(time.perf_counter()-t00)/N*1e6
print(rank, {k:round(v/N*1e6,1) for k,v in T.items()}, 'us; loop', round(tot,1), 'dev scan/dg/ev0-ev5 us', [round(float(x)*1e3,1) for x in np.mean(dev[10:],axis=0)], flush=True)
ex.close(); t.close(); dist.barrier(); dist.destroy_process_group()
