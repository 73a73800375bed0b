#!/usr/bin/env python3
"""Regenerate oracle/tables.json from the reference's thermodynamic tables.

TEST INFRASTRUCTURE ONLY (see oracle/README.md).

Reads the TAB-separated, '"'-quoted csv files the reference installs to
share/degen_primer (reference data/internal-NN.csv, data/dangling-ends.csv,
data/loops.csv, data/3-4-hairpin-loops.csv; parsed like DegenPrimer/UnifiedNN.py:163-199) and writes them
as one JSON document in this repo's own layout.  The JSON travels to the GPU
box; /root/reference does not.

usage: python oracle/make_tables.py [/root/reference/data]
"""
import csv
import json
import os
import sys


def read_table(path):
    with open(path, newline='') as fh:
        rows = list(csv.reader(fh, delimiter='\t', quotechar='"'))
    head = rows[0][1:]
    return {r[0]: {h: float(v) for h, v in zip(head, r[1:])} for r in rows[1:] if r}


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else '/root/reference/data'
    nn = read_table(os.path.join(src, 'internal-NN.csv'))
    ends = read_table(os.path.join(src, 'dangling-ends.csv'))
    loops = read_table(os.path.join(src, 'loops.csv'))
    tri_tetra = read_table(os.path.join(src, '3-4-hairpin-loops.csv'))
    doc = {
        'source': 'allista/DegenPrimer data/*.csv (SantaLucia & Hicks 2004)',
        'units': {'dH': 'kcal/mol', 'dS': 'cal/(K*mol)', 'dG': 'kcal/mol at 37C'},
        # [dH, dS, dG] triples keyed by the reference's "XY/ZW" strings
        'stacks': {k: [v['dH'], v['dS'], v['dG']] for k, v in nn.items()},
        'dangling': {k: [v['dH'], v['dS'], v['dG']] for k, v in ends.items()},
        # loop size -> [internal dG37, hairpin dG37]
        'loops': {k: [v['internal'], v['hairpin']] for k, v in loops.items()},
        # closing pair + loop letters of special tri-/tetra-loops -> [dH, dG37 bonus]
        'tri_tetra_hairpin_loops': {k: [v['dH'], v['dG']] for k, v in tri_tetra.items()},
    }
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'tables.json')
    with open(out, 'w') as fh:
        json.dump(doc, fh, indent=1, sort_keys=True)
    print('wrote', out, len(doc['stacks']), len(doc['dangling']), len(doc['loops']))


if __name__ == '__main__':
    main()
