"""Drop-in for the reference's ``swfd`` package surface (swfd/__init__.py:1-5)."""
from .frequent_directions import RobustFrequentDirections, FrequentDirectionsWithDump
from .seq_based_swfd import SeqBasedSWFD, TimeBasedSWFD, MultiStreamSWFD
from .lmfd import LMFD
from .difd import DIFD
from .fast_swfd import FastFrequentDirectionsWithDump, FastSlidingWindowFD, FastTimeBasedSWFD

__all__ = ["RobustFrequentDirections", "FrequentDirectionsWithDump", "SeqBasedSWFD", "TimeBasedSWFD", "MultiStreamSWFD", "LMFD", "DIFD",
           "FastFrequentDirectionsWithDump", "FastSlidingWindowFD", "FastTimeBasedSWFD"]
