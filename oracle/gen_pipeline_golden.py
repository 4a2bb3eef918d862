"""Record the UNMODIFIED reference pipeline (``src/pipeline.py``) calling its bundle adjuster on real frames.

    python -m oracle.gen_pipeline_golden       # from the repo root; needs /root/reference, cv2, scipy (~6 min)

TEST INFRASTRUCTURE.  The reference ships three desk photographs (``desk_images/``) and no video, so the
frames are fed back and forth (11, 22, 33, 22, 11, ...) through ``VisualOdometryPipeline.process_frame`` -
ORB, matching, pose estimation, key-frame selection and triangulation all run as written - and the run ends
with the global bundle adjustment of ``src/main.py:81-87`` (``window_size`` mutated to the number of
keyframes and restored).  ``pipeline.BundleAdjuster`` is replaced by a subclass of the reference's own
class that only RECORDS: the complete map before the call, the attribute ``window_size`` at the call, what
``run`` printed, and the poses / points afterwards.  ``tests/test_gpu_dropin.py`` rebuilds each recorded map
from the mirror classes and runs the B200 ``BundleAdjuster`` on it (caller integration).

The recording is stored as ``tests/golden/pipeline_desk.npz`` (keys ``c<k>_*`` for call k).
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle.reference_loader import load_reference  # noqa: E402

FRAME_ORDER = [11, 22, 33, 22, 11, 22, 33, 22]
KEYFRAME_CRITERIA = {"min_pixel_displacement": 20.0, "min_rotation": 0.15, "min_feature_ratio": 0.25,
                     "min_parallax_deg": 1.0, "min_tracked_for_parallax": 20}          # src/main.py:27-33
CAMERA_MATRIX = np.array([[912.7820434570312, 0.0, 650.2929077148438],
                          [0.0, 913.0294189453125, 362.7241516113281],
                          [0.0, 0.0, 1.0]])                                            # src/main.py:36-40


def dump_map(gmap) -> dict:
    """Everything ``BundleAdjuster.run`` can read from a map, as flat arrays."""
    kf_ids = sorted(gmap.keyframes)
    mp_ids = sorted(gmap.map_points)
    obs_kf, obs_mp, obs_kp, obs_uv = [], [], [], []
    for i in kf_ids:
        kf = gmap.keyframes[i]
        for mp_id, kp_idx in kf.observations:
            obs_kf.append(i); obs_mp.append(mp_id); obs_kp.append(kp_idx); obs_uv.append(kf.keypoints[kp_idx].pt)
    return dict(
        kf_ids=np.array(kf_ids, dtype=np.int64),
        kf_R=np.array([gmap.keyframes[i].R for i in kf_ids], dtype=np.float64).reshape(-1, 3, 3),
        kf_t=np.array([gmap.keyframes[i].t for i in kf_ids], dtype=np.float64).reshape(-1, 3, 1),
        kf_nkp=np.array([len(gmap.keyframes[i].keypoints) for i in kf_ids], dtype=np.int64),
        obs_kf=np.array(obs_kf, dtype=np.int64), obs_mp=np.array(obs_mp, dtype=np.int64),
        obs_kp=np.array(obs_kp, dtype=np.int64), obs_uv=np.array(obs_uv, dtype=np.float64).reshape(-1, 2),
        mp_ids=np.array(mp_ids, dtype=np.int64),
        mp_pos=np.array([gmap.map_points[i].position for i in mp_ids], dtype=np.float64).reshape(-1, 3, 1),
    )


def main():
    import cv2
    import scipy
    ref_ba, _ = load_reference()
    import features      # reference modules (flat imports, /root/reference/src on sys.path)
    import parameters
    import pipeline
    print(f"numpy {np.__version__} scipy {scipy.__version__} cv2 {cv2.__version__}")

    calls = []

    class RecordingBundleAdjuster(ref_ba.BundleAdjuster):
        def run(self, gmap):
            before = dump_map(gmap)
            out = io.StringIO()
            t0 = time.time()
            with contextlib.redirect_stdout(out):
                super().run(gmap)
            wall = time.time() - t0
            sys.stdout.write(out.getvalue())
            after = dump_map(gmap)
            calls.append(dict(window_size=int(self.window_size), before=before, after=after, wall=wall,
                              lines=[ln for ln in out.getvalue().splitlines() if "LBA" in ln]))

    pipeline.BundleAdjuster = RecordingBundleAdjuster      # the one injection point (src/pipeline.py:39)

    frames = {n: cv2.imread(f"/root/reference/desk_images/image{n}.png") for n in set(FRAME_ORDER)}
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.chdir(tmp)
        try:
            for d in [parameters.OUTPUT_DIR, parameters.TRAJECTORY_DIR_2D, parameters.TRAJECTORY_DIR_3D,
                      *parameters.DEBUG_DIRS.values()]:
                os.makedirs(d, exist_ok=True)
            vo = pipeline.VisualOdometryPipeline(camera_matrix=CAMERA_MATRIX, dist_coeffs=np.zeros(5),
                                                 feature_extractor=features.ORBExtractor(n_features=4000),
                                                 feature_matcher=features.BruteForceMatcher(),
                                                 keyframe_criteria=KEYFRAME_CRITERIA)
            for n in FRAME_ORDER:
                vo.process_frame(frames[n])
            # src/main.py:81-87
            num_keyframes = vo.map.next_keyframe_id
            if num_keyframes > 2:
                original = vo.bundle_adjuster.window_size
                vo.bundle_adjuster.window_size = num_keyframes
                vo.bundle_adjuster.run(vo.map)
                vo.bundle_adjuster.window_size = original
        finally:
            os.chdir(cwd)

    gold = dict(n_calls=np.int64(len(calls)), camera_matrix=CAMERA_MATRIX,
                ba_window_size=np.int64(parameters.BA_WINDOW_SIZE), frame_order=np.array(FRAME_ORDER, dtype=np.int64),
                versions=np.array([np.__version__, scipy.__version__, cv2.__version__]))
    for k, c in enumerate(calls):
        gold[f"c{k}_window_size"] = np.int64(c["window_size"])
        gold[f"c{k}_lines"] = np.array(c["lines"])
        gold[f"c{k}_ref_wall_s"] = np.float64(c["wall"])
        for name, v in c["before"].items():
            gold[f"c{k}_{name}"] = v
        for name in ("kf_R", "kf_t", "mp_pos"):
            gold[f"c{k}_out_{name}"] = c["after"][name]
        print(f"call {k}: window_size {c['window_size']} KFs {len(c['before']['kf_ids'])} MPs {len(c['before']['mp_ids'])} "
              f"obs {len(c['before']['obs_kf'])} ref {c['wall']:.1f} s | {c['lines'][-1].strip()}")
    path = os.path.join(ROOT, "tests", "golden", "pipeline_desk.npz")
    np.savez_compressed(path, **gold)
    print(f"-> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


if __name__ == "__main__":
    main()
