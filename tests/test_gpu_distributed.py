"""Two ranks on two GPUs (NCCL): the sharded fused call with the in-stream hit
exchange gives the whole-template hit set.  Needs >= 2 GPUs (gpurun --gpus 2);
skipped on a single-GPU box.  The host logic is covered on CPU with gloo in
tests/test_distributed_cpu.py."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRIMERS = ['ATATTCTACRACGGCTATCC', 'GAASGCRAAKATYGGGAAC']


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, text, k, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    os.environ['DEGENPRIMER_B200_DEVICE'] = str(rank)
    dist.init_process_group('nccl', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world,
                            device_id=torch.device('cuda', rank))
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import distributed as D
    from degenprimer_b200.SeqUtils import unambiguous_sequences
    tdf.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=55.0))
    ctx = tdf._ctx()
    sh = D.ShardedTemplate(ctx, text, rank, world, max_primer_len=max(map(len, PRIMERS)))
    q = ctx.prepare(PRIMERS, k, [unambiguous_sequences(p) for p in PRIMERS])
    ex = D.HitExchange(ctx, max_hits=512)
    hits, dimers, local, st, dst = sh.find(q, tdf.conditions(), exchange=ex)
    np.save(os.path.join(out_dir, 'fast%d.npy' % rank), hits)
    # the exchange keeps three calls
    for _ in range(3):
        ctx.find(sh.template, q, cond=tdf.conditions())
    older, newest = ex.collect_hits(1), ex.collect_hits(0)
    ctx.find_begin(sh.template, q, tdf.conditions())
    oldest = ex.collect_hits(2)          # on the host already: collected while the new call runs
    ctx.find_end()
    same = older is not None and newest is not None and np.array_equal(older, newest) and len(older) > 0
    same = same and oldest is not None and np.array_equal(oldest, newest)
    # the kernel-side conversion + sort of the gathered hits against the host-side one
    os.environ['DEGENPRIMER_B200_EXCHANGE_SORT'] = 'host'
    by_host = ex.collect_hits(0)
    del os.environ['DEGENPRIMER_B200_EXCHANGE_SORT']
    same = same and by_host is not None and np.array_equal(by_host, newest)
    np.save(os.path.join(out_dir, 'ages%d.npy' % rank), np.array([int(same)]))
    ex.close()
    small = D.HitExchange(ctx, max_hits=1)               # forces the fallback gather
    hits2, _, _, _, _ = sh.find(q, tdf.conditions(), exchange=small)
    small.close()
    np.save(os.path.join(out_dir, 'slow%d.npy' % rank), hits2)
    np.save(os.path.join(out_dir, 'ndim%d.npy' % rank), np.array([len(dimers), len(local)]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_gpu_sharded_find(tmp_path, ctx):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    from degenprimer_b200 import TD_Functions as tdf
    from degenprimer_b200 import distributed as D
    from tests.util import random_dna, plant_sites
    rng = np.random.default_rng(5)
    text = random_dna(rng, 400_000)
    text, _ = plant_sites(rng, text, PRIMERS, 2, 40)
    cut = D.segments(len(text), 2)[0][1]
    text = text[:cut - 7] + 'ATATTCTACAACGGCTATCC' + text[cut + 13:]      # a site straddling the cut
    k = 2
    whole = ctx.template_from_sequences([text])
    want, _ = ctx.scan(whole, PRIMERS, k)
    assert len(want) > 50
    mp.spawn(_worker, args=(2, _free_port(), text, k, str(tmp_path)), nprocs=2, join=True)
    n_local = 0
    for r in range(2):
        assert np.load(os.path.join(str(tmp_path), 'ages%d.npy' % r))[0] == 1
        for kind in ('fast', 'slow'):
            got = np.load(os.path.join(str(tmp_path), '%s%d.npy' % (kind, r)))
            assert np.array_equal(got['pos'], want['pos']) and np.array_equal(got['strand'], want['strand'])
            assert np.array_equal(got['primer'], want['primer']) and np.array_equal(got['score3'], want['score3'])
        n_local += int(np.load(os.path.join(str(tmp_path), 'ndim%d.npy' % r))[1])
    assert n_local == len(want)                     # every hit is owned by exactly one rank
