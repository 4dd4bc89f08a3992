"""Join an ncu SASS source page (per-instruction counters) with nvdisasm line info of the in-tree cubin, and print the
hottest source lines.  Usage: python tools/ncu_lines.py REPORT.ncu-rep [kernel-substring] [min-pct] [translation-unit]"""
import csv, re, subprocess, sys, tempfile, os, collections

rep = sys.argv[1]
ksub = sys.argv[2] if len(sys.argv) > 2 else "pose_kernel<(int)0, (bool)0>"
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
obj = sys.argv[4] if len(sys.argv) > 4 else "pose_kernels"  # translation unit the kernel lives in (csrc/<obj>.cu)
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ie, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = {h: i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h}
inst = [(r[1], float(r[ie] or 0), float(r[si] or 0), {h: float(r[i] or 0) for h, i in stall_cols.items()}) for r in rows[2:] if len(r) > ie]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", obj, os.environ.get("PERCH_LINES_LIB", os.path.join(root, "perception_b200/libperch_b200.so"))], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
# split per function
want_mode = re.search(r"\(int\)(\d), \(bool\)(\d)", ksub)
mangled = f"ILi{want_mode.group(1)}ELb{want_mode.group(2)}E" if want_mode else ksub
# every function whose mangled name matches; the profiled one is recognised by its instruction count (a substring such as
# "pose_kernel" matches all template instantiations: joining the wrong one misattributes every line)
funcs, cur, name = collections.OrderedDict(), None, None
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l) or re.match(r"\s*\.section\s+\.text\.(\S+),", l)
    if m:
        name = m.group(1).split(",")[0] if mangled in m.group(1) else None
        if name:
            funcs.setdefault(name, [])
        continue
    if not name:
        continue
    m = re.search(r"//## File \"([^\"]+)\", line (\d+)(.*)", l)
    if m:
        inl = re.search(r"inlined at \"([^\"]+)\", line (\d+)", m.group(3))
        cur = (os.path.basename(m.group(1)), int(m.group(2)), (os.path.basename(inl.group(1)), int(inl.group(2))) if inl else None)
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        funcs[name].append((cur, m.group(2)))
match = [n for n, v in funcs.items() if len(v) == len(inst)]
if len(match) != 1:
    sys.exit(f"error: no unique function with {len(inst)} instructions among {[(n[-40:], len(v)) for n, v in funcs.items()]}: "
             "the library on disk is not the build that was profiled (PERCH_LINES_LIB=path of the right one)")
lines = funcs[match[0]]
print(f"function {match[0]} ({len(lines)} instructions)")
tot_i = sum(a for _, a, _, _ in inst); tot_s = sum(s for _, _, s, _ in inst)
agg = collections.OrderedDict()
for (loc, _), (_, a, s, st) in zip(lines, inst):
    key = loc[2] if (loc and loc[2] and loc[2][0] == (obj + ".cu")) else (loc[0], loc[1]) if loc else ("?", 0)
    # attribute to outermost pose_kernels.cu line where possible
    if loc and loc[0] != (obj + ".cu") and loc[2]:
        key = loc[2]
    elif loc:
        key = (loc[0], loc[1])
    e = agg.setdefault(key, [0, 0, collections.Counter()])
    e[0] += a; e[1] += s
    for h, v in st.items():
        e[2][h] += v
src = open(os.path.join(root, "perception_b200/csrc/" + obj + ".cu")).read().splitlines()
print(f"total warp-instr {tot_i:.3e}, samples {tot_s:.0f}")
for key in sorted(agg, key=lambda k: (k[0], k[1])):
    a, s, st = agg[key]
    if 100 * a / tot_i >= minpct or 100 * s / tot_s >= minpct:
        top = ",".join(f"{h[6:]}:{100*v/max(s,1):.0f}" for h, v in st.most_common(3))
        text = src[key[1] - 1].strip()[:90] if key[0] == (obj + ".cu") and 0 < key[1] <= len(src) else ""
        print(f"{key[0]}:{key[1]:<5} inst {100*a/tot_i:5.1f}%  samp {100*s/tot_s:5.1f}%  [{top}]  {text}")

# ---- segments between barriers -------------------------------------------------------------------------------
print("\nsegments between BAR.SYNC (instruction stream order):")
seg_a = seg_s = 0; lo = 10**9; hi = 0; st_acc = collections.Counter(); nseg = 0
def flush():
    global seg_a, seg_s, lo, hi, st_acc, nseg
    top = ",".join(f"{h[6:]}:{100*v/max(seg_s,1):.0f}" for h, v in st_acc.most_common(4))
    print(f"  seg{nseg:<3} lines {lo}-{hi}: inst {100*seg_a/tot_i:5.1f}%  samp {100*seg_s/tot_s:5.1f}%  [{top}]")
    seg_a = seg_s = 0; lo = 10**9; hi = 0; st_acc = collections.Counter(); nseg += 1
for (loc, txt), (_, a, s, st) in zip(lines, inst):
    seg_a += a; seg_s += s
    for h, v in st.items(): st_acc[h] += v
    if loc and loc[0] == (obj + ".cu"):
        lo = min(lo, loc[1]); hi = max(hi, loc[1])
    if "BAR.SYNC" in txt:
        flush()
flush()
