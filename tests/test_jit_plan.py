"""The prefilter of the run-time specialised scan kernel (csrc/dp_jit.cu) is a necessary condition: every
window the ORACLE calls a hit must have at least `need` completely matching position groups.  The plan comes
from dp_jit_describe (host only, no GPU); the hits from oracle O2 on templates with planted near-hits."""
import numpy as np
import pytest

from degenprimer_b200 import _native as nat
from oracle import search_o2 as o2
from oracle.sequtils import reverse_complement
from tests.util import plant_sites, random_dna

SETS = {'A': 'A', 'C': 'C', 'G': 'G', 'T': 'T', 'R': 'AG', 'Y': 'CT', 'S': 'GC', 'W': 'AT', 'K': 'GT', 'M': 'AC',
        'B': 'CGT', 'D': 'AGT', 'H': 'ACT', 'V': 'ACG', 'N': 'ACGT'}

CASES = [
    (['ACTTGTCATGGCGACTGTCC', 'CCATGAGCGCTGCTCAGATA'], 1),
    (['AGRGTTYGATYMTGGCTCWG', 'TACGGYTACCTTGTTACGASTT'], 2),
    (['AGRGTTYGATYMTGGYTYWG', 'KACGGYTAMCTTRTTAYGASTW'], 3),
    (['GAASGCRAAKATYGGGAAC'], 0),
    (['ACGGTCATTGCAGTCCGATAGGCTTAAGCCTGATC'], 4),
    (['ACGTBDHVNNACGTRYSWKMACGTACG', 'TTGACCAGTNACGTACGGTCA'], 2),
]


def _alive(window, master, strand, groups, need):
    """The group count of the plan on one forward-strand window."""
    seen = reverse_complement(window) if strand else window          # what the primer is compared with
    L = len(master)
    ok = 0
    for g in groups:
        full = True
        for j in g:                                                   # j = position in the WINDOW
            pj = L - 1 - j if strand else j
            if seen[pj] not in SETS[master[pj]]:
                full = False
        ok += full
    return ok >= need


@pytest.mark.parametrize('case', range(len(CASES)))
def test_every_oracle_hit_passes_the_prefilter(case):
    masters, k = CASES[case]
    plans, mode = nat.jit_plan(masters, k)
    assert len(plans) == 2 * len(masters) and mode in ('shared-planes', 'folded-planes')
    rng = np.random.default_rng(100 + case)
    text = random_dna(rng, 60000)
    text, _ = plant_sites(rng, text, masters, max(k, 1), 40)
    checked = 0
    for plan in plans:
        pi, strand = plan['pattern'] // 2, plan['pattern'] % 2
        m = masters[pi]
        L = len(m)
        assert plan['K'] == L - max(1, L - k) and plan['need'] == len(plan['groups']) - plan['K'] >= 1
        used = [j for g in plan['groups'] for j in g]
        assert len(used) == len(set(used)) and all(0 <= j < L for j in used)      # disjoint groups inside the window
        res = o2.find(text, m, k)
        hits = res[5] if strand else res[4]
        for pos in hits:
            fwd = len(text) - L - int(pos) if strand else int(pos)
            assert _alive(text[fwd:fwd + L], m, strand, plan['groups'], plan['need']), (plan, fwd)
            checked += 1
        assert plan['p_alive'] < 2e-3
    assert checked >= 10


def test_multiplex_plan_shares_planes():
    from degenprimer_b200.synthetic import multiplex_primers
    plans, mode = nat.jit_plan(multiplex_primers(), 3)
    assert mode == 'shared-planes' and len(plans) == 40
    assert max(len(p['groups']) for p in plans) <= 9
