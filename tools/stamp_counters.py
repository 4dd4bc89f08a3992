"""Write profiles/kernel_counters.json and profiles/dram_traffic.json from metrics-only ncu passes of the fused kernel
(one CSV per workload, produced on the GPU box by
  ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
      --clock-control none -k regex:pose_kernel -c 1 --csv --log-file gpurun_out/<prefix>_<w>.csv python tools/ab_cost.py --child <w>)
and stamp them with the hash of the kernel sources they were captured from (bench.kernel_source_hash): bench.py uses the
counters only while that hash matches.  Usage: python tools/stamp_counters.py gpurun_out/r2_final 'note'"""
import csv
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402

prefix, note = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
counters, traffic = {}, {}
for w in ("c1", "c2", "c3", "c5"):
    f = Path(f"{prefix}_{w}.csv")
    if not f.exists():
        continue
    vals, grid = {}, None
    for r in csv.reader(f.read_text().splitlines()):
        if len(r) > 14 and r[0].isdigit():
            vals[r[12]] = float(r[14].replace(",", ""))
            grid = int(r[8].strip("()").split(",")[0])
    counters[w] = {"warp_inst_per_launch": int(vals["smsp__inst_executed.sum"]), "poses": grid, "sm_count": 148,
                   "kernel_ns_under_ncu": vals["gpu__time_duration.sum"],
                   "source": f"ncu --metrics smsp__inst_executed.sum, one launch of pose_kernel<MODE_COST> ({f.name}){'; ' + note if note else ''}",
                   "source_sha16": bench.kernel_source_hash()}
    traffic[w] = int(vals["dram__bytes_read.sum"] + vals.get("dram__bytes_write.sum", 0))
(ROOT / "profiles/kernel_counters.json").write_text(json.dumps(counters, indent=1))
traffic["source"] = "dram__bytes_read.sum + dram__bytes_write.sum of one launch (same ncu passes as kernel_counters.json)"
(ROOT / "profiles/dram_traffic.json").write_text(json.dumps(traffic, indent=1))
print(json.dumps(counters, indent=1))
