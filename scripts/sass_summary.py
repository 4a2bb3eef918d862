"""Per-kernel SASS mnemonic histogram of libba_b200.so (cuobjdump -sass), the evidence for the kernel choices:
FP64 arithmetic (DFMA/DADD/DMUL), bulk asynchronous copies (UBLKCP = cp.async.bulk, SYNCS = mbarrier), no
floating-point atomics (ATOM/RED absent), shuffles, barriers.

    python scripts/sass_summary.py [out_dir] [tag]   -> <out_dir>/<tag>_sass_summary.txt, <tag>_sass_tile_pass.txt  (tag: r02t)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles")
tag = sys.argv[2] if len(sys.argv) > 2 else "r02t"
so = os.path.join(ROOT, "bundle_adjustment_b200", "libba_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", so], capture_output=True, text=True, check=True).stdout
usage = {}
for m in re.finditer(r"Function (\S+):\n\s*(.*)", res):
    usage[m.group(1)] = m.group(2).strip()
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
WATCH = ["DFMA", "DADD", "DMUL", "MUFU", "LDG", "STG", "LDS", "STS", "LDGSTS", "UBLKCP", "SYNCS", "BAR", "SHFL", "ATOM", "ATOMS", "RED",
         "REDUX", "CCTL", "ACQBULK", "UCGABAR_ARV", "UCGABAR_WAIT", "LDL", "STL", "ERRBAR", "MEMBAR"]
funcs = re.split(r"\n\s*Function : ", sass)[1:]
lines = [f"SASS of {os.path.basename(so)} (sm_100a), cuobjdump -sass; columns = instruction counts per kernel", ""]
tile = None
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    ops = collections.Counter()
    for ln in f.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m:
            ops[m.group(1)] += 1
    total = sum(ops.values())
    dn = demangle(name)
    if "cub::" in dn.split("(")[0]:   # library sort / scan kernels of the one-time problem construction
        continue
    if "k_tile_pass<(int)0, (bool)1>" in dn:
        tile = (dn, f)
    lines.append(f"{dn}")
    lines.append(f"    {usage.get(name, '')}")
    lines.append(f"    instructions {total}: " + "  ".join(f"{k} {ops[k]}" for k in WATCH if ops[k]))
    fp_atomics = [ln.strip() for ln in f.splitlines() if re.search(r"\b(ATOM|RED|ATOMG|ATOMS)\b.*\.(F64|F32)", ln)]
    lines.append(f"    floating-point atomics: {len(fp_atomics)}")
    lines.append("")
os.makedirs(out_dir, exist_ok=True)
with open(os.path.join(out_dir, tag + "_sass_summary.txt"), "w") as fh:
    fh.write("\n".join(lines))
if tile:
    with open(os.path.join(out_dir, tag + "_sass_tile_pass.txt"), "w") as fh:
        body = tile[1].split("\n", 1)[1]
        keep = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln).rstrip() for ln in body.splitlines()]
        fh.write(f"Function : {tile[0]}\n" + "\n".join(ln for ln in keep if ln.strip()) + "\n")
print("\n".join(lines[:60]))
