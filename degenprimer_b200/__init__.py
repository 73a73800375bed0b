"""degenprimer_b200 — the data-parallel hot path of allista/DegenPrimer on B200.

Scan (annealing-site search), duplex dG / per-variant Tm and the template
loader, as hand-written sm_100a CUDA behind a C-ABI (include/degenprimer_b200.h),
under the reference's own Python call signatures:

    from degenprimer_b200.SearchEngine import SearchEngine
    from degenprimer_b200.Primer import Primer
    from degenprimer_b200.SeqView import SeqView
    from degenprimer_b200 import TD_Functions as tdf

Importing the package touches neither the shared library nor CUDA; the first
call that needs the GPU loads libdegenprimer_b200.so and raises if it (or a
usable sm_100 device) is missing.  There is no CPU fallback.
"""
__version__ = '0.1.0'
