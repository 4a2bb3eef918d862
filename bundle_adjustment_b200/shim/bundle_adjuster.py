"""Flat-import shim: with THIS directory ahead of the reference's ``src/`` on ``sys.path``, the reference's own
``from bundle_adjuster import BundleAdjuster`` (``src/pipeline.py:12``, ``src/main.py``) resolves to the B200 drop-in, and
nothing else of the reference changes.  (The package itself uses relative imports and cannot be imported flat.)"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from bundle_adjustment_b200.bundle_adjuster import BundleAdjuster  # noqa: E402,F401
