"""Import the UNMODIFIED reference (``/root/reference/src``) in the build container.  TEST INFRASTRUCTURE.

The reference imports ``open3d`` and ``matplotlib`` at module scope (``src/bundle_adjuster.py:6``,
``src/map_structures.py:4``, ``src/visualization.py:4-8``); neither is installed and both are only used
for debug artefacts (sparsity PNG, ``.pcd`` dump), so they are replaced by ``MagicMock`` modules.  No
reference source is copied or altered.  ``/root/reference`` does not exist on the GPU box: nothing that
runs there may call :func:`load_reference` - it is used only by ``oracle/gen_golden.py`` (to produce the
committed fixtures) and by CPU tests that skip when the path is absent.
"""
from __future__ import annotations

import importlib
import os
import sys
from unittest import mock

REFERENCE_SRC = "/root/reference/src"


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_SRC, "bundle_adjuster.py"))


def load_reference():
    """-> (bundle_adjuster module, map_structures module) of the reference."""
    if not reference_available():
        raise RuntimeError("reference sources are not present on this machine")
    for name in ("open3d", "matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d"):
        if name not in sys.modules:
            sys.modules[name] = mock.MagicMock(name=name)
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    ba = importlib.import_module("bundle_adjuster")
    ms = importlib.import_module("map_structures")
    if not ba.__file__.startswith(REFERENCE_SRC):
        raise RuntimeError(f"'bundle_adjuster' resolved to {ba.__file__}, not the reference")
    return ba, ms
