"""profiles/traffic.json from an ncu summary (scripts/ncu_summary.py output): dram read + write bytes per launch of each hot
kernel, keyed by the bench's kernel names.

    python scripts/make_traffic.py profiles/r02_ncu_hot_kernels.csv venice
"""
import csv
import json
import os
import sys

NAMES = [("k_tile_pass<0", "schur_matvec_tiles"), ("k_tile_pass<1", "backsub_tiles"), ("k_tile_lin", "lin_points"),
         ("k_pcg_tail", "pcg_tail"), ("k_camera_pass<0", "lin_cameras"), ("k_camera_pass<1", "schur_rhs_cameras"),
         ("k_point_prepare", "point_prepare")]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
src, shape = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(src)))
hdr, units, data = rows[0], rows[1], rows[2:]
ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
out = {}
for r in data:
    for pat, name in NAMES:
        if pat.replace("<", "<(int)") in r[ik] or pat in r[ik]:
            b = float(r[ir]) * UNIT[units[ir]] + float(r[iw]) * UNIT[units[iw]]
            out.setdefault(name, []).append(b)
            break
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
doc = json.load(open(path)) if os.path.isfile(path) else {}
doc[shape] = {k: round(max(v)) for k, v in out.items()}   # the largest capture of a kind (the first launch of a kind can be a short one)
doc["_source"] = f"dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full capture summarised in {src} ({shape} shape, 1 GPU)"
json.dump(doc, open(path, "w"), indent=1)
print(json.dumps(doc[shape], indent=1))
