"""Run the bench workloads (fused COST kernel, lattice CLOUD stage, forced bands / tiny tiles) through the bounds-checked
build of the per-pose kernel (`make check` -> perception_b200/libperch_b200_check.so): every data-dependent shared-memory
index is range checked on the device and a violation fails the call.  compute-sanitizer (memcheck / racecheck) is closed
on this pool; this is the substitute.  Usage (GPU box): python tools/bounds_check.py > profiles/r2_bounds_check.md"""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from perception_b200 import _capi  # noqa: E402

lib = _capi.load_library(ROOT / "perception_b200" / "libperch_b200_check.so")
_capi._lib = lib
import bench  # noqa: E402

print("# Bounds-checked build of pose_kernel (make check): all workloads\n")
print("| workload | poses | tile cap | stage | status | checksum |\n|---|---|---|---|---|---|")
for name in ("c1", "c2", "c3", "c5"):
    r, wl = bench.build_workload(name, 0, 0, None if name != "c5" else 20000)
    kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"],
              color_distance_threshold=20.0, occlusion_threshold=1.0)
    for cap in (None, 64):  # 64 lattice cells per band: many bands, queue and deferral lists under pressure
        if cap:
            os.environ["PERCH_COST_TILE_CAP"] = str(cap)
        else:
            os.environ.pop("PERCH_COST_TILE_CAP", None)
        out = r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)
        print(f"| {name} | {wl['P']} | {cap or 2048} | COST | ok | {float(np.nansum(out['rendered_cost'])):.3f} |")
    os.environ.pop("PERCH_COST_TILE_CAP", None)
    n = min(wl["P"], 256)
    out = r.render_cuda_multi_unified("CLOUD", wl["poses"][:n], wl["pmm"][:n], wl["plabel"][:n] if wl["plabel"] is not None else None,
                                      stride=wl["stride"])
    print(f"| {name} | {n} | 6144 | CLOUD | ok | {out['result_cloud_point_num']} points |")
    if wl["cost_type"] != 1:
        kw1 = dict(kw, cost_type=1)
        out = r.render_cuda_multi_unified("COST", wl["poses"][:n], wl["pmm"][:n], None, wl["tot"][:n], **kw1)
        print(f"| {name} | {n} | 2048 | COST colour | ok | {float(np.nansum(out['rendered_cost'])):.3f} |")
    r.close()
print("\nNo call returned the bounds-check error (PERCH_ERR_CUDA, flag 256).")
