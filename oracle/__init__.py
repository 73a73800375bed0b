"""CPU restatement of the DegenPrimer hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/--impl reference
legs may import this package (see oracle/README.md).  Parity status: pinned to
outputs of the reference itself (its Python-2.7 sources run through oracle/refshim.py,
recorded in tests/golden/reference_golden.json.gz); the reference ships no golden vectors.
"""
