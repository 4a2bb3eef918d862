import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with `-m gpu` on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """Make sure libba_b200.so exists (nvcc cross-compiles without a GPU; ~30 s the first time)."""
    from bundle_adjustment_b200 import build
    build.build()


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz"))


def load_golden(name):
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    spec = {k[5:]: g[k] for k in g.files if k.startswith("spec_")}
    return g, spec


def mirror_map(spec):
    """Rebuild the fixture's map from the repo's own map classes (no reference, no cv2)."""
    from bundle_adjustment_b200 import synthetic as S
    from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint, Map, MapPoint
    return S.build_map(spec, Map, Keyframe, MapPoint, KeyPoint.from_xy)


def oracle_problem(spec, flat):
    from oracle import ba_oracle as O
    K = spec["camera_matrix"]
    return O.Problem(K[0, 0], K[1, 1], K[0, 2], K[1, 2], len(flat["adj_kf_ids"]), len(flat["mp_ids"]),
                     flat["fixed_rvec"], flat["fixed_tvec"], flat["obs_cam"], flat["obs_pt"], flat["obs_uv"])


@pytest.fixture(scope="session")
def gpu_solver():
    """One device handle for the whole GPU session.  Fails (does not skip) if the library or the
    device is missing: GPU tests must never pass on a fallback."""
    from bundle_adjustment_b200 import FlatSolver
    s = FlatSolver(0)
    yield s
    s.close()
