"""distributed.GenomeLayout — the strong-scaling split of a multi-record genome: pieces,
halo, owned window starts, translation of piece hits into whole-genome coordinates.  CPU: the per-rank
scan is stood in by the oracle run on that rank's pieces (world-size-2 gloo for the gather); on a GPU box
the same layout drives dp_find (tests/test_gpu_round2.py, bench.py --gpus N)."""
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


def _genome(seed):
    from tests.util import random_dna, plant_sites
    rng = np.random.default_rng(seed)
    recs = []
    for n in (5003, 40, 0, 9001, 64, 2500, 17):
        t = random_dna(rng, n)
        if n > 1000:
            t, _ = plant_sites(rng, t, PRIMERS, 2, 6)
        recs.append(t)
    return recs


def _oracle_hits(texts, primers, k, owned_end=None):
    """Hits of a multi-record template as dp_find reports them (record = index in `texts`),
    window starts restricted to the first `owned_end` padded positions."""
    from degenprimer_b200._native import HIT_DTYPE
    from oracle import search_o2 as o2
    rows, start = [], 0
    for ri, text in enumerate(texts):
        for pi, m in enumerate(primers):
            if len(text) < len(m):
                continue
            r = o2.find(text, m, k)
            L = len(m)
            for p in r[4]:
                if owned_end is None or start + int(p) < owned_end:
                    rows.append((int(p), ri, pi, 0, 0))
            for p in r[5]:
                fwd = len(text) - L - int(p)
                if owned_end is None or start + fwd < owned_end:
                    rows.append((int(p), ri, pi, 1, 0))
        start += (len(text) + 31) // 32 * 32
    return np.array(rows, dtype=HIT_DTYPE) if rows else np.zeros(0, dtype=HIT_DTYPE)


def _rank_hits(layout, recs, rank, k):
    from degenprimer_b200 import distributed as D
    lens = [len(p) for p in PRIMERS]
    texts = [recs[r][off:off + n] for r, off, n, _own in layout.pieces[rank]]
    local = _oracle_hits(texts, PRIMERS, k, owned_end=layout.seg_end(rank))
    return layout.to_global(local, rank, lens)


def test_layout_covers_every_window_start_once():
    from degenprimer_b200.distributed import GenomeLayout
    lens = [5003, 40, 0, 9001, 64, 2500, 17]
    for world in (1, 2, 3, 5, 8, 16):
        lay = GenomeLayout(lens, world, 24)
        owned = {}
        for g in range(world):
            starts = lay.local_starts(g)
            for j, (r, off, n, own) in enumerate(lay.pieces[g]):
                assert off % 32 == 0 and 0 < own <= n and off + n <= lens[r] and starts[j] % 32 == 0
                assert n == min(lens[r] - off, own + 23)                       # halo = L_max - 1, never past the record
                owned.setdefault(r, []).append((off, off + own))
            if lay.pieces[g]:
                assert all(p[3] == p[2] for p in lay.pieces[g][:-1])            # only the last piece carries a halo
        for r, n in enumerate(lens):
            spans = sorted(owned.get(r, []))
            assert ([a for a, _ in spans] == [0] + [b for _, b in spans[:-1]] and spans[-1][1] == n) if n else not spans
        base, ls, off, rec, rl = lay.exchange_tables()
        assert len(base) == world + 1 and base[-1] == len(ls) == sum(len(p) for p in lay.pieces)
        assert [lens[r] for r in rec] == rl


@pytest.mark.parametrize('world', [2, 3, 7])
def test_sharded_genome_equals_whole_with_the_oracle(world):
    from degenprimer_b200 import distributed as D
    recs = _genome(world)
    k = 2
    lay = D.GenomeLayout([len(r) for r in recs], world, max(len(p) for p in PRIMERS))
    parts = [_rank_hits(lay, recs, g, k) for g in range(world)]
    got = D.sort_hits(np.concatenate(parts))
    want = D.sort_hits(_oracle_hits(recs, PRIMERS, k))
    assert len(want) > 20 and np.array_equal(got, want)


def _worker(rank, world, port, seed, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from degenprimer_b200 import distributed as D
    dist.init_process_group('gloo', init_method='tcp://127.0.0.1:%d' % port, rank=rank, world_size=world)
    recs = _genome(seed)
    lay = D.GenomeLayout([len(r) for r in recs], world, max(len(p) for p in PRIMERS))
    mine = _rank_hits(lay, recs, rank, 2)
    allh = D.sort_hits(D.all_gather_records(mine))
    np.save(os.path.join(out_dir, 'rank%d.npy' % rank), allh)
    dist.destroy_process_group()


def test_sharded_genome_gather_gloo(tmp_path):
    """world_size 2 over gloo: every rank ends with the hit set of the whole genome."""
    from degenprimer_b200 import distributed as D
    world, seed = 2, 11
    mp.spawn(_worker, args=(world, _free_port(), seed, str(tmp_path)), nprocs=world, join=True)
    want = D.sort_hits(_oracle_hits(_genome(seed), PRIMERS, 2))
    for rank in range(world):
        got = np.load(os.path.join(str(tmp_path), 'rank%d.npy' % rank))
        assert np.array_equal(got, want)
