"""cuda_icp package API: batched perch_icp_point2plane against the reference's own ICP_Point2Plane_cuda<Scene_projective>
(oracle/_ref/libicp_ref.so, one cloud per call as its API has it), 30 forced iterations, 400-point clouds on a 640x480 scene.
The reference harness re-uploads the scene on every call; its loop time is taken as t(30 iterations) - t(0 iterations).
Usage (GPU box): python tools/icp_probe.py [n_clouds]"""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import icp_cases  # noqa: E402
from oracle_lib import ICP_REF_SO, ReferenceIcp  # noqa: E402
from perception_b200.renderer import Renderer  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
case, K = icp_cases.scene()
s = case.scene
with Renderer(0) as r:
    pcd, nrm = r.icp_set_scene(s.source_depth, icp_cases.W, icp_cases.H, K, want_buffers=True)
    clouds, _ = icp_cases.clouds(pcd, s.source_label, n_clouds=B)
    row = {"clouds": B, "points_per_cloud": int(np.mean([len(c) for c in clouds])), "iterations": 30}
    for nn in (None, 0.01):
        wall, dev = [], []
        for _ in range(4):
            t0 = time.perf_counter()
            out = r.icp_point2plane(clouds, relative_fitness=0.0, relative_rmse=0.0, max_iteration=30, nn_max_dist=nn)
            wall.append(time.perf_counter() - t0)
            dev.append(r.last_timing()[0])
        tag = "projective" if nn is None else "nn"
        row[f"perch_{tag}_wall_ms"] = 1e3 * min(wall[1:])
        row[f"perch_{tag}_kernel_ms"] = float(np.median(dev[1:]))
        row[f"perch_{tag}_clouds_per_s"] = B / min(wall[1:])
        row[f"perch_{tag}_mean_fitness"] = float(np.mean(out["fitness"]))
    if ICP_REF_SO.exists():
        ref = ReferenceIcp()
        n_ref = min(B, 16)
        t_full, t_zero, fit = [], [], []
        for c in clouds[:n_ref]:
            for _ in range(2):
                t0 = time.perf_counter()
                o = ref.point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, relative_fitness=0.0, relative_rmse=0.0, max_iteration=30)
                t1 = time.perf_counter()
                ref.point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, relative_fitness=0.0, relative_rmse=0.0, max_iteration=0)
                t2 = time.perf_counter()
            t_full.append(t1 - t0); t_zero.append(t2 - t1); fit.append(o["fitness"])
        loop = float(np.median(np.array(t_full) - np.array(t_zero)))
        row.update({"reference_loop_ms_per_cloud": 1e3 * loop, "reference_call_ms_per_cloud": 1e3 * float(np.median(t_full)),
                    "reference_clouds_per_s_loop_only": 1.0 / loop, "reference_mean_fitness": float(np.mean(fit))})
    print(json.dumps(row))
