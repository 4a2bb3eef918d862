"""Tier T0 on config 1: the UNMODIFIED reference ``BundleAdjuster.run`` (imported from /root/reference, build container
only, CPU) on the same synthetic pipeline window scripts/run_latency.py gives to the drop-in.  Wall time of one ``run``."""
import io
import os
import sys
import tempfile
import time
from contextlib import redirect_stdout

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cv2  # noqa: E402
from bundle_adjustment_b200 import synthetic as S  # noqa: E402
from oracle.reference_loader import load_reference  # noqa: E402

ba, ms = load_reference()
n_kfs = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n_pts = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
spec = S.make_window_spec(n_kfs=n_kfs, n_pts=n_pts)
gmap = S.build_map(spec, ms.Map, ms.Keyframe, ms.MapPoint, lambda x, y: cv2.KeyPoint(float(x), float(y), 1.0))
adj = ba.BundleAdjuster(spec["camera_matrix"], window_size=5)
cwd = os.getcwd()
with tempfile.TemporaryDirectory() as d:
    os.chdir(d)
    os.makedirs("debug_sparsity", exist_ok=True); os.makedirs("output_map/lba_steps", exist_ok=True)
    t0 = time.perf_counter()
    with redirect_stdout(io.StringIO()) as out:
        adj.run(gmap)
    dt = time.perf_counter() - t0
    os.chdir(cwd)
print(f"reference BundleAdjuster.run on the {n_kfs}-keyframe / {n_pts}-point window: {dt:.1f} s on one core ({os.cpu_count()} present)")
print(out.getvalue().strip().splitlines()[-1])
