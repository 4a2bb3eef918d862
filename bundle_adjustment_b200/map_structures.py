"""Host-side map data model consumed and mutated by the B200 bundle adjuster.

Mirrors the attribute layout of the reference's ``src/map_structures.py:7-54`` so that a
``Map`` built by the reference pipeline (``src/pipeline.py``) and one built from these
classes are interchangeable at the ``BundleAdjuster.run(gmap)`` boundary:

* ``Keyframe.R`` 3x3 float64 (world -> camera), ``Keyframe.t`` 3x1, ``Keyframe.keypoints``
  a list of objects with a ``.pt`` 2-tuple (``cv2.KeyPoint`` in the pipeline),
  ``Keyframe.observations`` a list of ``(map_point_id, keypoint_index)``.
* ``MapPoint.position`` 3x1 float64, ``MapPoint.observations`` a list of
  ``(keyframe_id, keypoint_index)``.
* ``Map.keyframes`` / ``Map.map_points`` dicts id -> object; ``add_*`` raise ``ValueError``
  on a duplicate id (reference ``src/map_structures.py:44-45, 51-52``).

The Open3D point-cloud export (reference ``get_pcd``, ``:56-78``) needs a graphics-side
package that is unavailable on this box; ``get_points_colors`` returns the same arrays as
plain numpy instead.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Tuple

import numpy as np


@dataclass
class KeyPoint:
    """Minimal stand-in for ``cv2.KeyPoint``: only ``.pt`` is read by bundle adjustment
    (reference ``src/bundle_adjuster.py:216``).  Coordinates are float32-valued, as OpenCV's are."""
    pt: Tuple[float, float]

    @staticmethod
    def from_xy(x: float, y: float) -> "KeyPoint":
        return KeyPoint((float(np.float32(x)), float(np.float32(y))))


@dataclass
class MapPoint:
    id: int
    position: np.ndarray
    observations: List[Tuple[int, int]] = field(default_factory=list)
    color: Any = None


@dataclass
class Keyframe:
    id: int
    R: np.ndarray
    t: np.ndarray
    keypoints: list = field(default_factory=list)
    descriptors: Any = None
    observations: List[Tuple[int, int]] = field(default_factory=list)
    img: Any = None


class Map:
    def __init__(self) -> None:
        self.keyframes: Dict[int, Keyframe] = {}
        self.map_points: Dict[int, MapPoint] = {}
        self.next_keyframe_id = 0
        self.next_map_point_id = 0

    def add_keyframe(self, keyframe: Keyframe) -> None:
        if keyframe.id in self.keyframes:
            raise ValueError(f"Keyframe with ID {keyframe.id} already exists.")
        self.keyframes[keyframe.id] = keyframe
        self.next_keyframe_id += 1

    def add_map_point(self, map_point: MapPoint) -> None:
        if map_point.id in self.map_points:
            raise ValueError(f"MapPoint with ID {map_point.id} already exists.")
        self.map_points[map_point.id] = map_point
        self.next_map_point_id += 1

    def get_points_colors(self):
        """(N,3) positions and (N,3) colours of all map points, float64 (what the reference
        hands to Open3D in ``get_pcd``)."""
        if not self.map_points:
            return np.zeros((0, 3)), np.zeros((0, 3))
        pos = np.array([np.asarray(p.position, dtype=np.float64).reshape(3) for p in self.map_points.values()])
        col = np.array([np.zeros(3) if p.color is None else np.asarray(p.color, dtype=np.float64).reshape(3)
                        for p in self.map_points.values()])
        return pos, col
