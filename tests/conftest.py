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
    """Window fixtures (one reference ``run`` each); recordings of whole pipelines are named pipeline_*."""
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and not f.startswith(("pipeline_", "converged_")))


def recorded_map(g, k, which="before"):
    """Map of call k of a pipeline recording (``oracle/gen_pipeline_golden.py``), rebuilt from the repo's
    own map classes.  which = "before" (what ``run`` was given) or "after" (what the reference left)."""
    from bundle_adjustment_b200.map_structures import Keyframe, KeyPoint, Map, MapPoint
    pre = f"c{k}_"
    out = "" if which == "before" else "out_"
    gmap = Map()
    kf_ids = g[pre + "kf_ids"].tolist()
    obs_kf, obs_mp, obs_kp, obs_uv = (g[pre + n] for n in ("obs_kf", "obs_mp", "obs_kp", "obs_uv"))
    for i, kid in enumerate(kf_ids):
        rows = np.nonzero(obs_kf == kid)[0]
        kps = [None] * int(g[pre + "kf_nkp"][i])
        for r in rows:
            kps[int(obs_kp[r])] = KeyPoint((float(obs_uv[r, 0]), float(obs_uv[r, 1])))
        gmap.add_keyframe(Keyframe(id=kid, R=g[pre + out + "kf_R"][i].copy(), t=g[pre + out + "kf_t"][i].copy(), keypoints=kps,
                                   observations=[(int(obs_mp[r]), int(obs_kp[r])) for r in rows]))
    for i, mid in enumerate(g[pre + "mp_ids"].tolist()):
        gmap.add_map_point(MapPoint(id=mid, position=g[pre + out + "mp_pos"][i].copy()))
    return gmap


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
