"""GICP against a dense scene cloud (BASELINE configs[3]): kernel time of 4096 poses x 20 linearisations, and a check that
the pruned grid search returns what the streamed linear scan returns (PERCH_GICP_STREAM_MAX large = every range streamed).
Usage (GPU box): python tools/gicp_dense_probe.py"""
import os, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import bench
from perception_b200 import synth
r, wl = bench.build_workload("c2", 0, 0, 4096)
s = wl["scene"]
cam = (synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
dense = r.depth2cloud_global(s.source_depth, s.source_color, 640, 480, 1, *cam, 1)
r.set_observed(dense["result_cloud_eigen"], dense["result_cloud_color"], None)
args = dict(stride=8, sensor_resolution=0.01, occlusion_threshold=1.0, icp_k_correspondences=10, icp_max_iterations=20,
            icp_rotation_epsilon=0.0, icp_transformation_epsilon=0.0)
os.environ["PERCH_GICP_STALL_WINDOW"] = "0"
g = r.gicp(wl["poses"], wl["pmm"], None, **args)
ms = []
for _ in range(3):
    g = r.gicp(wl["poses"], wl["pmm"], None, **args); ms.append(r.last_timing()[0])
print("n_obs", dense["rendered_cloud_point_num"], "kernel ms", np.median(ms), "iterations", g["iterations"].mean())
# warm-start radius (in fine cells) beyond which a search starts cold: PERCH_GICP_WARM_CELLS
default_wc = os.environ.get("PERCH_GICP_WARM_CELLS")
for wc in ([] if os.environ.get("PERCH_PROBE_NO_SWEEP") else (0, 3, 6, 12, 24, 48, 100000)):
    os.environ["PERCH_GICP_WARM_CELLS"] = str(wc)
    ms2 = []
    for _ in range(3):
        g2 = r.gicp(wl["poses"], wl["pmm"], None, **args); ms2.append(r.last_timing()[0])
    print("warm cells", wc, "kernel ms", np.median(ms2), "same transforms", np.array_equal(g2["transforms"], g["transforms"]))
if default_wc is None:
    os.environ.pop("PERCH_GICP_WARM_CELLS", None)
else:
    os.environ["PERCH_GICP_WARM_CELLS"] = default_wc
for kap in ([] if not os.environ.get("PERCH_PROBE_KAPPA_SWEEP") else (2, 5, 10, 20, 40, 80, 160, 400)):
    os.environ["PERCH_GICP_LEVEL_KAPPA_X10"] = str(kap)
    ms2 = []
    for _ in range(3):
        g2 = r.gicp(wl["poses"], wl["pmm"], None, **args); ms2.append(r.last_timing()[0])
    print("kappa x10", kap, "kernel ms", np.median(ms2), "same transforms", np.array_equal(g2["transforms"], g["transforms"]))
os.environ.pop("PERCH_GICP_LEVEL_KAPPA_X10", None)
# points per cell (x10) of the finer grid GICP builds for dense clouds: PERCH_GICP_GRID_TARGET_X10 (read when the grid is built)
for tgt in ([] if not os.environ.get("PERCH_PROBE_GRID_SWEEP") else (10, 20, 30, 50, 80, 120)):
    os.environ["PERCH_GICP_GRID_TARGET_X10"] = str(tgt)
    r.set_observed(dense["result_cloud_eigen"], dense["result_cloud_color"], None)
    ms2 = []
    for _ in range(4):
        g2 = r.gicp(wl["poses"], wl["pmm"], None, **args); ms2.append(r.last_timing()[0])
    print("grid target x10", tgt, "kernel ms", np.median(ms2[1:]), "same transforms", np.array_equal(g2["transforms"], g["transforms"]))
os.environ.pop("PERCH_GICP_GRID_TARGET_X10", None)
r.set_observed(dense["result_cloud_eigen"], dense["result_cloud_color"], None)
n = 256
os.environ["PERCH_GICP_STREAM_MAX"] = str(1 << 30)
t0 = time.time()
h = r.gicp(wl["poses"][:n], wl["pmm"][:n], None, **args)
print("streamed reference for", n, "poses:", time.time() - t0, "s; transforms equal:", np.array_equal(h["transforms"], g["transforms"][:n]),
      "max |d|", np.abs(h["transforms"] - g["transforms"][:n]).max())
