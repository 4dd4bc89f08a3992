"""Dump the profiled SASS of one kernel with per-instruction executed counts (per-pose warp-instr) and line info.
Usage: python tools/ncu_sass.py REPORT.ncu-rep MODE COLOR n_poses > out.txt"""
import csv, re, subprocess, sys, tempfile, os
rep, mode, color, npose = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ie, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
inst = [(r[1], float(r[ie] or 0), float(r[si] or 0)) for r in rows[2:] if len(r) > ie]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "pose_kernels", os.path.join(root, "perception_b200/libperch_b200.so")], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
mangled = f"ILi{mode}ELb{color}E"
lines, cur, on = [], None, False
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l) or re.match(r"\s*\.section\s+\.text\.(\S+),", l)
    if m:
        on = mangled in m.group(1); continue
    if not on: continue
    m = re.search(r"//## File \"([^\"]+)\", line (\d+)", l)
    if m:
        cur = f"{os.path.basename(m.group(1))}:{m.group(2)}"; continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: lines.append((cur, m.group(2)))
    elif re.match(r"\s*\.L_x_\d+:", l): lines.append((None, l.strip()))
j = 0
for loc, txt in lines:
    if loc is None:
        print(txt); continue
    _, a, s = inst[j]; j += 1
    print(f"{a/npose:9.1f} {s:6.0f}  {loc:24s} {txt}")
