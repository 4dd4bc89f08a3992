"""Device time of the RENDER stage (pose_kernel<MODE_IMAGE>: full-resolution depth + RGB images, 7 B/px/pose written)
against the HBM write roofline.  Usage (GPU box): python tools/render_probe.py [n_poses]"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
r, wl = bench.build_workload("c2", 0, 0, None)
poses, pmm, lab = wl["poses"][:n], wl["pmm"][:n], wl["plabel"][:n]
ms = []
for rep in range(4):
    out = r.render_cuda_multi_unified("RENDER", poses, pmm, lab)
    ms.append(r.last_timing()[0])
k = float(np.median(ms[1:]))
bytes_out = n * bench.W * bench.H * 7
peak, kind = bench.measured_peak_gbs()
print(json.dumps({"stage": "RENDER", "poses": n, "kernel_ms": k, "poses_per_s": n / (k / 1e3),
                  "written_GB_per_s": bytes_out / (k / 1e3) / 1e9, "hbm_peak_GB_per_s": peak, "frac": bytes_out / (k / 1e3) / 1e9 / peak,
                  "covered_px_per_pose": float((out["result_depth"] > 0).sum() / n)}))
