"""BASELINE configs[3] probe: GICP of the top poses against the FULL-resolution scene cloud (stride 1, ~300k points).
Usage (GPU box): python tools/c4_probe.py [n_top] [render_stride]"""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
from perception_b200 import synth  # noqa: E402

n_top = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
rs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
r, wl = bench.build_workload("c2", 0, 0, None)
kw = dict(stride=wl["stride"], cost_type=2, calculate_observed_cost=True, sensor_resolution=wl["res"], occlusion_threshold=1.0)
out = r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **kw)
cost = np.where(out["rendered_cost"] >= 0, out["rendered_cost"] + np.nan_to_num(out["observed_cost"], nan=1e9), 1e9)
sel = np.sort(np.argsort(cost, kind="stable")[:n_top])
s = wl["scene"]
cam = (synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
t0 = time.perf_counter()
obs = r.depth2cloud_global(s.source_depth, s.source_color, bench.W, bench.H, 1, *cam, 1)
print("scene cloud points:", obs["rendered_cloud_point_num"], f"depth2cloud {time.perf_counter() - t0:.3f} s")
t0 = time.perf_counter()
r.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], None)
print(f"set_observed {time.perf_counter() - t0:.3f} s")
args = dict(stride=rs, sensor_resolution=wl["res"], occlusion_threshold=1.0, icp_max_iterations=20, icp_k_correspondences=10,
            icp_rotation_epsilon=0.0, icp_transformation_epsilon=0.0)
for rep in range(3):
    t0 = time.perf_counter()
    g = r.gicp(wl["poses"][sel], wl["pmm"][sel], None, **args)
    w = time.perf_counter() - t0
    print(f"rep {rep}: perch_gicp wall {w * 1e3:.1f} ms, kernel {r.last_timing()[0]:.2f} ms, iterations {g['iterations'].mean():.1f}, "
          f"finite {np.isfinite(g['transforms']).all()}")
t = g["transforms"]
print("mean |t| (m):", float(np.linalg.norm(t[:, 9:], axis=1).mean()), " poses/s (wall):", len(sel) / w)
