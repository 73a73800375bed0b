"""world_size-2 gloo test of the N>1 host path: segment bounds, halo, coordinate
translation, ownership and the variable-length hit gather.  The per-rank scan is
stood in by the oracle run on that rank's slice (no GPU here); on a GPU box the
same code path calls dp_scan (tests/test_gpu_distributed.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PRIMERS = ['ATATTCTACRACGGCTATCC', 'GAASGCRAAKATYGGGAAC', 'ACGTTGCAAGCTTGCATGCAAGGT']


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_hits(text, primers, k):
    from degenprimer_b200._native import HIT_DTYPE
    from oracle import search_o2 as o2
    rows = []
    for pi, m in enumerate(primers):
        if len(text) < len(m):
            continue
        r = o2.find(text, m, k)
        rows += [(int(p), 0, pi, 0, 0) for p in r[4]] + [(int(p), 0, pi, 1, 0) for p in r[5]]
    return np.array(rows, dtype=HIT_DTYPE) if rows else np.zeros(0, dtype=HIT_DTYPE)


def _worker(rank, world, port, text, k, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from degenprimer_b200 import distributed as D
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    lens = [len(p) for p in PRIMERS]
    bb, be, sb, se = D.shard_bounds(len(text), world, rank, max(lens))
    local = _oracle_hits(text[bb:be], PRIMERS, k)
    hits = D.localize_hits(local, bb, be - bb, len(text), lens)
    hits = D.keep_owned(hits, sb, se, len(text), lens)
    allh = D.sort_hits(D.all_gather_records(hits))
    np.save(os.path.join(out_dir, 'rank%d.npy' % rank), allh)
    empty = D.all_gather_records(hits[:0])
    assert len(empty) == 0
    dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_sharded_scan_equals_whole(tmp_path, world):
    from degenprimer_b200 import distributed as D
    from tests.util import random_dna, plant_sites
    rng = np.random.default_rng(world)
    text = random_dna(rng, 20011)
    text, _ = plant_sites(rng, text, PRIMERS, 2, 14)
    # sites straddling the cut between rank 0 and rank 1
    cut = D.segments(len(text), world)[0][1]
    text = text[:cut - 7] + 'ATATTCTACAACGGCTATCC' + text[cut + 13:]
    text = text[:cut - 23] + 'GTTCCCGATTTTCGCCTTC' + text[cut - 4:]
    k = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, text, k, str(tmp_path)), nprocs=world, join=True)
    want = D.sort_hits(_oracle_hits(text, PRIMERS, k))
    assert len(want) > 20
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), 'rank%d.npy' % r))
        assert np.array_equal(got, want)


def test_segments_cover_and_align():
    from degenprimer_b200 import distributed as D
    for T in (1, 31, 32, 33, 1000, 83257441):
        for w in (1, 2, 3, 8):
            seg = D.segments(T, w)
            assert seg[0][0] == 0 and seg[-1][1] == T
            assert all(a[1] == b[0] for a, b in zip(seg, seg[1:]))
            assert all(b % 32 == 0 for b, _ in seg if b < T)


def _pairs_worker(rank, world, port, seqs, pairs, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from degenprimer_b200 import distributed as D
    from oracle import conditions as ocond
    from oracle.structures import find_dimers
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    ocond.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.3e-3, DNA=1e-10, DMSO=0.0, PCR_T=40.0))
    b, e = D.shard_pairs(len(pairs), world, rank)
    rec = np.zeros(e - b, dtype=[('pair', '<u4'), ('n', '<u4'), ('min_dG', '<f8')])
    for k, (i, j) in enumerate(pairs[b:e]):
        dimers = find_dimers(seqs[i], seqs[j])[0]
        rec[k] = (b + k, len(dimers), min((d.dG for d in dimers), default=0.0))
    allr = D.all_gather_records(rec)
    np.save(os.path.join(out_dir, 'pairs%d.npy' % rank), allr[np.argsort(allr['pair'])])
    dist.destroy_process_group()


def test_sharded_dimer_pairs_equal_whole(tmp_path):
    """N3 shards by sequence pair: two ranks, each its slice (the oracle stands in for dp_find_dimers), one
    variable-length gather; every rank ends up with the records of all pairs."""
    from degenprimer_b200 import distributed as D
    from oracle import conditions as ocond
    from oracle.structures import find_dimers
    for n, w in ((7, 2), (8, 3), (2, 4), (0, 2)):
        cuts = [D.shard_pairs(n, w, r) for r in range(w)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n and all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
        assert max(e - b for b, e in cuts) - min(e - b for b, e in cuts) <= 1
    rng = np.random.default_rng(3)
    seqs = [''.join(rng.choice(list('ACGT'), size=int(n))) for n in rng.integers(12, 30, size=9)]
    seqs += ['GGGAAATTTCCCGGGAAATTTCCC', 'GCGGCCGCTTTTGCGGCCGC']
    pairs = [(i, j) for i in range(len(seqs)) for j in range(i, len(seqs))][:41]
    port = _free_port()
    mp.spawn(_pairs_worker, args=(2, port, seqs, pairs, str(tmp_path)), nprocs=2, join=True)
    ocond.PCR_P.set(dict(Na=50e-3, Mg=3e-3, dNTP=0.3e-3, DNA=1e-10, DMSO=0.0, PCR_T=40.0))
    want = [(k, len(find_dimers(seqs[i], seqs[j])[0])) for k, (i, j) in enumerate(pairs)]
    for rank in range(2):
        got = np.load(os.path.join(str(tmp_path), 'pairs%d.npy' % rank))
        assert [(int(r['pair']), int(r['n'])) for r in got] == want
    assert sum(n for _, n in want) > 5
