"""A/B of two builds of libperch_b200.so on the COST stage of the bench workloads (device time of the fused kernel).
Each library runs in its own process (PERCH_B200_LIB); symbols an older build lacks are skipped.
Usage (GPU box): python tools/ab_cost.py perception_b200/_ab/libperch_old.so [more .so ...] [workloads...]"""
import ctypes as C
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def child(workloads):
    from perception_b200 import _capi
    probe = C.CDLL(os.environ.get("PERCH_B200_LIB", str(_capi.LIB_PATH)))
    _capi.SIGNATURES = {k: v for k, v in _capi.SIGNATURES.items() if hasattr(probe, k)}
    import numpy as np
    import bench
    for w in workloads:
        r, wl = bench.build_workload(w, 0, 0, None)
        kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"])
        ms = []
        for _ in range(8):
            out = r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)
            ms.append(r.last_timing()[0])
        print(json.dumps({"lib": os.environ.get("PERCH_B200_LIB", "current"), "workload": w, "poses": len(wl["poses"]),
                          "kernel_ms_median": float(np.median(ms[2:])), "kernel_ms_min": float(min(ms[2:])),
                          "cost_checksum": float(np.nansum(out["rendered_cost"]))}), flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "--child":
        child(sys.argv[2:])
    else:
        libs = [a for a in sys.argv[1:] if a.endswith(".so")]
        workloads = [a for a in sys.argv[1:] if not a.endswith(".so")] or ["c1", "c3", "c2"]
        for lib in (*libs, None, *libs, None):
            env = dict(os.environ)
            if lib:
                env["PERCH_B200_LIB"] = lib
            else:
                env.pop("PERCH_B200_LIB", None)
            subprocess.run([sys.executable, __file__, "--child", *workloads], env=env, check=True)
