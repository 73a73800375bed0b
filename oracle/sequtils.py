"""ORACLE (test infrastructure, see oracle/README.md) — IUPAC helpers.

Restates DegenPrimer/SeqUtils.py:30-60 (expansion table and order) and
DegenPrimer/SecStructures.py:34-55 (complement map, reverse_complement).
"""

# DegenPrimer/SeqUtils.py:30-40 — the ORDER of each list defines variant order.
EXPANSIONS = {
    'R': 'AG', 'Y': 'CT', 'S': 'GC', 'W': 'AT', 'K': 'GT', 'M': 'AC',
    'B': 'CGT', 'D': 'AGT', 'H': 'ACT', 'V': 'ACG', 'N': 'ATCG',
}
PLAIN = 'ATGC'

# DegenPrimer/SecStructures.py:34-48
COMPLEMENT = dict(zip('ATGCRYSWKMBVDHN', 'TACGYRSWMKVBHDN'))


def reverse_complement(seq):
    """DegenPrimer/SecStructures.py:50-55 — upper-cases, KeyError on foreign bytes."""
    return ''.join(COMPLEMENT[c] for c in seq.upper())[::-1]


def unambiguous_sequences(s):
    """DegenPrimer/SeqUtils.py:44-60.

    The growing list is tiled once per replacement and block r receives
    replacement r, so the EARLIEST degenerate position varies fastest.
    """
    out = ['']
    for letter in s:
        n = len(out)
        repl = EXPANSIONS.get(letter)
        if repl is None:
            out = [x + letter for x in out]
        else:
            out = [out[i % n] + repl[i // n] for i in range(n * len(repl))]
    return out
