"""Seeded synthetic templates for the benchmark configs.

There is no network and no chromosome download: templates are i.i.d. uniform
ACGT from numpy's PCG64 with primer sites planted at seeded positions (random
templates have ~no hits at small k), optionally with long runs of N.  The
TP53 / 16S primer strings are builder-chosen: the reference names the files
(`README.md:75-76`) but does not ship them.
"""
import numpy as np

from .SeqUtils import reverse_complement, unambiguous_sequences

_ACGT = np.frombuffer(b'ACGT', dtype=np.uint8)

# builder-fixed primers -----------------------------------------------------------
TP53_F = 'ACTTGTCATGGCGACTGTCC'
TP53_R = 'CCATGAGCGCTGCTCAGATA'
# 16S 27F / 1492R-like with IUPAC degeneracy: 2*2*2*2*2 = 32 and 2*2*2 = 8 variants
# (includes W and S so the 2/3-weight letters are exercised)
P27F = 'AGRGTTYGATYMTGGCTCWG'
P1492R = 'TACGGYTACCTTGTTACGASTT'

# north-star pair: 2 x 128 = 256 unambiguous variants (R Y M K S W: full- and 2/3-weight letters)
P27F_D128 = 'AGRGTTYGATYMTGGYTYWG'
P1492R_D128 = 'KACGGYTAMCTTRTTAYGASTW'

# 20-plex (config 4): L = 18..25, four degenerate primers (V = 2, 8, 16, 64)

def multiplex_primers():
    rng = np.random.default_rng(np.random.PCG64(4242))
    out = []
    for i in range(20):
        L = 18 + (i % 8)
        seq = ''.join('ACGT'[x] for x in rng.integers(0, 4, size=L))
        out.append(seq)
    # degenerate ones: substitute two-fold letters (incl. S / W) at seeded positions
    two_fold = {'A': 'RWM', 'C': 'YSM', 'G': 'RSK', 'T': 'YWK'}
    for idx, n_deg in ((3, 1), (7, 3), (11, 4), (15, 6)):
        seq = list(out[idx])
        for pos in rng.choice(len(seq), size=n_deg, replace=False):
            seq[pos] = two_fold[seq[pos]][int(rng.integers(3))]
        out[idx] = ''.join(seq)
    return out


SWEEP_35 = 'ACGGTCATTGCAGTCCGATAGGCTTAAGCCTGATC'          # config 5: prefixes of a fixed 35-mer
SWEEP_35_V16 = 'ACGGTCRTTGCAGTYCGATAGGMTTAAGCKTGATC'      # same with four two-fold letters (V = 16)

WORKLOADS = {
    # name: (total bp, n_records, seed, primers, k, n_plant pairs, N-run bp)
    'config1_tp53_5Mbp': dict(bp=5_000_000, records=1, seed=1, primers=[TP53_F, TP53_R], k=1,
                              plant=32, n_runs=0),
    'config2_tp53_chr17': dict(bp=83_257_441, records=1, seed=2, primers=[TP53_F, TP53_R], k=1,
                               plant=64, n_runs=2, n_run_len=3_000_000),
    'config3_16S_500Mbp': dict(bp=500_000_000, records=200, seed=3, primers=[P27F, P1492R], k=2,
                               plant=4, n_runs=0),
    'config3_16S_256var_500Mbp': dict(bp=500_000_000, records=200, seed=3, primers=[P27F_D128, P1492R_D128], k=2,
                                      plant=4, n_runs=0),
    'config4_20plex_3Gbp': dict(bp=3_100_000_000, records=24, seed=4, primers=None, k=3, plant=40, n_runs=0,
                                plant_all=True),
    'northstar_256fold_3Gbp': dict(bp=3_000_000_000, records=24, seed=6, primers=[P27F_D128, P1492R_D128], k=3,
                                   plant=40, n_runs=2, n_run_len=2_000_000),
    'config5_sweep_1Gbp': dict(bp=1_000_000_000, records=1, seed=5, primers=[SWEEP_35], k=2, plant=64, n_runs=0),
}


def random_template(rng, n):
    """uint8 ASCII array of n uniform ACGT bases."""
    out = np.empty(n, dtype=np.uint8)
    step = 1 << 26
    for a in range(0, n, step):
        b = min(n, a + step)
        out[a:b] = _ACGT[rng.integers(0, 4, size=b - a, dtype=np.uint8)]
    return out


def _site(rng, master, k):
    variants = unambiguous_sequences(master)
    s = list(variants[int(rng.integers(len(variants)))])
    L = len(s)
    nsub = int(rng.integers(0, k + 1))
    for p in rng.choice(L, size=nsub, replace=False):
        s[p] = 'ACGT'[(('ACGT'.index(s[p])) + int(rng.integers(1, 4))) % 4]
    return ''.join(s)


def plant_pairs(rng, arr, primers, k, n_pairs, lo=0, hi=None, all_primers=False):
    """Plant n_pairs (forward site of primers[0], reverse-complement site of
    primers[-1] 100-5000 bp downstream) so that amplicons exist; with all_primers
    the pair is (primers[i], primers[i+1]) for a seeded i.  Returns the list of
    (primer index, strand, position)."""
    hi = len(arr) if hi is None else hi
    planted = []
    span = hi - lo
    if span < 6000:
        return planted
    for _ in range(n_pairs):
        at = lo + int(rng.integers(0, span - 5200))
        gap = int(rng.integers(100, 5000))
        fi, ri = 0, len(primers) - 1
        if all_primers:
            fi = int(rng.integers(len(primers)))
            ri = (fi + 1) % len(primers)
        f = _site(rng, primers[fi], k)
        r = reverse_complement(_site(rng, primers[ri], k))
        arr[at:at + len(f)] = np.frombuffer(f.encode(), dtype=np.uint8)
        arr[at + gap:at + gap + len(r)] = np.frombuffer(r.encode(), dtype=np.uint8)
        planted.append((fi, 0, at))
        planted.append((ri, 1, at + gap))
    return planted


def make_workload(name, scale=1.0, seed_offset=0):
    """Returns dict(records=[uint8 arrays], ids, primers, k, planted, bp)."""
    w = dict(WORKLOADS[name])
    if w['primers'] is None:
        w['primers'] = multiplex_primers()
    rng = np.random.default_rng(np.random.PCG64(w['seed'] + seed_offset))
    total = max(10_000, int(w['bp'] * scale))
    nrec = w['records']
    if nrec == 1:
        lens = [total]
    else:
        raw = np.exp(rng.uniform(np.log(0.5e6), np.log(8e6), size=nrec))     # log-uniform record lengths
        lens = np.maximum(1000, (raw / raw.sum() * total).astype(np.int64)).tolist()
    records, planted = [], []
    for i, n in enumerate(lens):
        arr = random_template(rng, int(n))
        pl = plant_pairs(rng, arr, w['primers'], w['k'], w['plant'], all_primers=bool(w.get('plant_all')))
        planted.append(pl)
        if w.get('n_runs') and n > 4 * w['n_run_len'] * scale + 10:
            run = max(100, int(w['n_run_len'] * scale))
            for j in range(w['n_runs']):
                at = int((j + 1) * n / (w['n_runs'] + 1))
                arr[at:at + run] = ord('N')
        records.append(arr)
    return dict(name=name, records=records, ids=['rec%03d' % i for i in range(nrec)],
                primers=list(w['primers']), k=w['k'], planted=planted, bp=int(sum(lens)))
