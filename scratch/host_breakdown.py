"""Host-side cost of one call, split into the C calls and their Python wrappers (2 ranks)."""
import os, sys, time, ctypes as C; sys.path.insert(0, '.')
import numpy as np, torch, torch.distributed as dist
rank=int(os.environ['RANK']); world=int(os.environ['WORLD_SIZE']); local=int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
os.environ['DEGENPRIMER_B200_DEVICE']=str(local)
from degenprimer_b200 import TD_Functions as tdf, _native as nat
from degenprimer_b200.synthetic import make_workload
from degenprimer_b200.SeqUtils import unambiguous_sequences
from degenprimer_b200.distributed import HitExchange
wl = make_workload('config2_tp53_chr17', seed_offset=1000*rank)
tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=62.0))
ctx = tdf._ctx(); cond = tdf.conditions()
t = ctx.template_from_sequences([wl['records'][0].tobytes()])
var=[unambiguous_sequences(p) for p in wl['primers']]
KK=int(os.environ.get('KK','1'))
q = ctx.prepare(wl['primers'], KK, var)
ex = HitExchange(ctx, max_hits=1024)
lib=ctx.lib
N=300
T=dict(begin_c=0.0, collect_c=0.0, collect_py=0.0, end_c=0.0, end_py=0.0)
hits, n_hits, dimers, n_dimers = C.c_void_p(), C.c_uint64(), C.c_void_p(), C.c_uint64()
st, dst = nat.dp_scan_stats(), nat.dp_dg_stats()
gh, gn = C.c_void_p(), C.c_uint64()
for it in range(N+10):
    if it==10:
        for k in T: T[k]=0.0
    a=time.perf_counter()
    lib.dp_find_begin(ctx._h, t._h, q.n, q.m, q.k, 0, 2**64-1, 3, 0, q.vc_p, q.vo_p, q.vb_p, C.byref(cond))
    b=time.perf_counter()
    if it>1:
        rc=lib.dp_comm_hits(ctx._h, 2, C.byref(gh), C.byref(gn))
        c=time.perf_counter()
        arr=ctx._take(gh, gn.value, nat.HIT_DTYPE)
        d=time.perf_counter()
    else:
        c=d=b
    torch.cuda.synchronize()      # GPU done: what remains in find_end is host work
    e=time.perf_counter()
    lib.dp_find_end(ctx._h, C.byref(hits), C.byref(n_hits), C.byref(dimers), C.byref(n_dimers), C.byref(st), C.byref(dst))
    f=time.perf_counter()
    h=ctx._take(hits, n_hits.value, nat.HIT_DTYPE); dd=ctx._take(dimers, n_dimers.value, nat.DIMER_DTYPE)
    g=time.perf_counter()
    for k,v in zip(T,(b-a,c-b,d-c,f-e,g-f)): T[k]+=v
print(rank, {k:round(v/N*1e6,1) for k,v in T.items()}, 'n gathered', gn.value, flush=True)
ex.close(); t.close(); dist.barrier(); dist.destroy_process_group()
