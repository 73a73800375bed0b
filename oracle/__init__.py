"""CPU restatement of the DegenPrimer hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/--impl reference
legs may import this package (see oracle/README.md).  Parity status: unpinned —
the reference ships no golden vectors for this path and cannot run in this image.
"""
