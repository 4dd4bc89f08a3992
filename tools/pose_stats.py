"""Per-level survivor counts of the fused kernel's candidate test on a bench workload (debug build: `make stats`).
Usage (GPU box): python tools/pose_stats.py [c2|c1]"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from perception_b200 import _capi  # noqa: E402

lib = _capi.load_library(ROOT / "perception_b200" / "libperch_b200_stats.so")
_capi._lib = lib  # every Renderer of this process uses the debug build
import bench  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
r, wl = bench.build_workload(name, 0, 0, None)
P = wl["P"]
kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"],
          color_distance_threshold=20.0, occlusion_threshold=1.0)
lib.perch_debug_stats.argtypes = [C.c_void_p, C.c_int]
lib.perch_debug_stats(None, 1)
out = r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)
st = (C.c_ulonglong * 16)()
lib.perch_debug_stats(st, 0)
s = np.array(list(st), np.float64) / P
names = ["quads sphere-tested (L1)", "L1 survivors (quads)", "quads through L2+L3", "L2+L3 survivors (triangles)",
         "exact: triangles with lattice cells", "exact: lattice cells visited", "exact: fragments passing",
         "cooperative triangles", "-", "L2 warp steps", "pass A/B rounds", "bands", "band cells"]
print(f"workload {name}: P={P}, per pose:")
for n, v in zip(names, s):
    print(f"  {n:40s} {v:10.1f}")
print("  rendered points/pose", float(np.mean(out["counts"][0])) if "counts" in out else None)
