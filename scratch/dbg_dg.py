import sys; sys.path.insert(0, '.')
import numpy as np
from tests.test_gpu_native import _oracle_dimers
from tests.util import random_dna, plant_sites
from degenprimer_b200 import TD_Functions as tdf
from degenprimer_b200.SeqUtils import unambiguous_sequences
master, k = 'GAASGCRAAKATYGGGAAC', 3
rng = np.random.default_rng(5)
t = random_dna(rng, 20000, 'ACGTN', p=[0.2495] * 4 + [0.002])
t, _ = plant_sites(rng, t, [master], k, 20)
cond = dict(Na=50e-3, Mg=3e-3, dNTP=0.35e-3, DNA=1e-9, PCR_T=55.0)
p, m, ref = _oracle_dimers(t, master, k, cond)
tdf.PCR_P.set(cond)
ctx = tdf._ctx()
tmpl = ctx.template_from_sequences([t])
hits, _ = ctx.scan(tmpl, [master], k)
variants = unambiguous_sequences(master)
L = len(master)
dimers, stats = ctx.duplex_dg(tmpl, hits, [L], [len(variants)], [0], ''.join(variants), tdf.conditions())
print(stats)
print(hits[:3])
for d in dimers[:24]:
    print(d['hit'], d['variant'], bin(d['match_mask']), d['dG'])
pos, dups = ref[0][0]
print('oracle pos', pos)
for d in dups[:3]:
    print(d.name, d.fwd_seq)
    for dm in d.dimers:
        print('   ', bin(sum(1 << i for i in dm.fwd_matches)), dm.dG)
