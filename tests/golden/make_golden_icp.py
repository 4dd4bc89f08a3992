"""Generate tests/golden/icp.npz from the reference's OWN cuda_icp sources (oracle/_ref/libicp_ref.so): scene preparation
(host code) and ICP_Point2Plane_cuda<Scene_projective> (GPU) on tests/icp_cases.py.
Needs a GPU: gpurun -- 'python tests/golden/make_golden_icp.py gpurun_out/golden'"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import icp_cases  # noqa: E402
from oracle_lib import Oracle, ReferenceIcp  # noqa: E402


def main(out_dir):
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    ref, orc = ReferenceIcp(), Oracle()
    case, K = icp_cases.scene()
    s = case.scene
    pcd, nrm = ref.scene(s.source_depth, icp_cases.W, icp_cases.H, K)
    o_pcd, o_nrm = orc.icp_scene(s.source_depth, icp_cases.W, icp_cases.H, K)
    print("scene: oracle == reference host code:", np.array_equal(pcd, o_pcd), np.array_equal(nrm, o_nrm))
    cl, _ = icp_cases.clouds(pcd, s.source_label)
    import hashlib
    save = {"scene_sha256": np.frombuffer(hashlib.sha256(pcd.tobytes() + nrm.tobytes()).digest(), np.uint8)}
    for i, c in enumerate(cl):
        runs = [ref.point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K) for _ in range(2)]
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K)
        r = runs[0]
        print(f"cloud {i}: n={len(c)} ref fitness {r['fitness']:.5f} rmse {r['inlier_rmse']:.6f} | reproducible "
              f"{np.array_equal(r['transformation'], runs[1]['transformation'])} | oracle dT "
              f"{np.abs(o['transformation'] - r['transformation']).max():.2e} dfit {abs(o['fitness'] - r['fitness']):.2e} "
              f"drmse {abs(o['inlier_rmse'] - r['inlier_rmse']):.2e}", flush=True)
        save[f"T{i}"], save[f"fit{i}"], save[f"rmse{i}"], save[f"cloud{i}"] = r["transformation"], np.float32(r["fitness"]), np.float32(r["inlier_rmse"]), r["cloud"]
        r3 = ref.point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_iteration=3)   # three updates: always comparable
        save[f"T3_{i}"], save[f"fit3_{i}"], save[f"rmse3_{i}"] = r3["transformation"], np.float32(r3["fitness"]), np.float32(r3["inlier_rmse"])
    np.savez_compressed(out / "icp.npz", **save)
    print("wrote", out / "icp.npz")
    # Scene_nn (kd-tree association, pcd_scene.h): the reference builds scene cloud, normals and tree itself
    save = {}
    for i, c in enumerate(cl):
        r0 = ref.point2plane_nn(c, s.source_depth, icp_cases.W, icp_cases.H, K, max_iteration=0)   # association only
        r = ref.point2plane_nn(c, s.source_depth, icp_cases.W, icp_cases.H, K)
        r2 = ref.point2plane_nn(c, s.source_depth, icp_cases.W, icp_cases.H, K)
        o0 = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_dist_diff=0.01, max_iteration=0, nn=True)
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_dist_diff=0.01, nn=True)
        print(f"nn cloud {i}: ref fitness0 {r0['fitness']:.5f} rmse0 {r0['inlier_rmse']:.6f} | oracle {o0['fitness']:.5f} "
              f"{o0['inlier_rmse']:.6f} | final ref {r['fitness']:.5f} {r['inlier_rmse']:.6f} oracle {o['fitness']:.5f} "
              f"{o['inlier_rmse']:.6f} dT {np.abs(o['transformation'] - r['transformation']).max():.2e} reproducible "
              f"{np.array_equal(r['transformation'], r2['transformation'])} it {o['iterations']}", flush=True)
        save[f"T{i}"], save[f"fit{i}"], save[f"rmse{i}"] = r["transformation"], np.float32(r["fitness"]), np.float32(r["inlier_rmse"])
        save[f"fit0_{i}"], save[f"rmse0_{i}"] = np.float32(r0["fitness"]), np.float32(r0["inlier_rmse"])
    np.savez_compressed(out / "icp_nn.npz", **save)
    print("wrote", out / "icp_nn.npz")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).parent))
