"""Drop-in for the reference's ``swfd`` package surface (swfd/__init__.py:1-5)."""
from .frequent_directions import RobustFrequentDirections, FrequentDirectionsWithDump
from .seq_based_swfd import SeqBasedSWFD, TimeBasedSWFD, MultiStreamSWFD

__all__ = ["RobustFrequentDirections", "FrequentDirectionsWithDump", "SeqBasedSWFD", "TimeBasedSWFD", "MultiStreamSWFD"]
