"""B200-native sparse bundle adjustment: drop-in for the local-BA step of egirgin/bundle_adjustment.

``BundleAdjuster`` mirrors the reference's class (``src/bundle_adjuster.py``); ``FlatSolver`` exposes
the same device solver on flat arrays; ``map_structures`` mirrors the reference's map data model.
The arithmetic lives in ``libba_b200.so`` (hand-written sm_100a CUDA, C ABI in ``include/ba_b200.h``).
"""
from .bundle_adjuster import BundleAdjuster
from .map_structures import Keyframe, KeyPoint, Map, MapPoint
from .solver import FlatSolver, make_options

__all__ = ["BundleAdjuster", "FlatSolver", "make_options", "Map", "Keyframe", "MapPoint", "KeyPoint"]
