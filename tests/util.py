"""Shared helpers for the tests: seeded synthetic templates with planted sites."""
import numpy as np

from oracle.sequtils import reverse_complement, unambiguous_sequences


def random_dna(rng, n, alphabet='ACGT', p=None):
    idx = rng.choice(len(alphabet), size=n, p=p)
    return np.frombuffer(alphabet.encode(), dtype=np.uint8)[idx].tobytes().decode()


def mutate(rng, s, n_sub, positions=None):
    s = list(s)
    if positions is None:
        positions = rng.choice(len(s), size=n_sub, replace=False)
    for p in positions:
        s[p] = rng.choice([b for b in 'ACGT' if b != s[p]])
    return ''.join(s)


def plant_sites(rng, template, masters, k, n_plant):
    """Overwrite random places with variants of the primers (fwd) / their reverse
    complements (rev), carrying 0..k substitutions incl. terminal ones."""
    t = list(template)
    planted = []
    T = len(t)
    for master in masters:
        variants = unambiguous_sequences(master)
        L = len(master)
        for i in range(n_plant):
            v = variants[int(rng.integers(len(variants)))]
            nsub = int(rng.integers(0, k + 1))
            pos_choice = None
            if nsub and i % 3 == 0:
                pos_choice = [L - 1] + list(rng.choice(L - 1, size=nsub - 1, replace=False))
            elif nsub and i % 3 == 1:
                pos_choice = [0] + list(rng.choice(np.arange(1, L), size=nsub - 1, replace=False))
            site = mutate(rng, v, nsub, pos_choice)
            strand = i % 2
            if strand:
                site = reverse_complement(site)
            at = int(rng.integers(0, T - L))
            t[at:at + L] = site
            planted.append((master, strand, at))
    return ''.join(t), planted
