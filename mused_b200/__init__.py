"""mused_b200 -- B200-native SWFD sketch-and-assign hot path of kelaendi/mused.

``mused_b200.swfd`` mirrors the reference's ``swfd`` package surface
(SeqBasedSWFD, FrequentDirectionsWithDump, RobustFrequentDirections) and
``mused_b200.matrix_operations`` the numeric operators main.py calls; all
arithmetic runs in hand-written sm_100a CUDA behind the C ABI in
``include/mused_b200.h``.  No CPU fallback.
"""
__version__ = "0.1.0"
