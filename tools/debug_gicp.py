import sys
sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import numpy as np
from oracle_lib import Oracle
from scenes import make_case, observed_3dof
from test_gicp import _perturbed, _angle_deg
from perception_b200 import synth
from perception_b200.renderer import Renderer
orc = Oracle()
case = make_case(2000, 3, 42); s = case.scene; stride = 4
obs = observed_3dof(case, stride, orc)
poses, pmm, lab = _perturbed(case, 9)
with Renderer(0) as r:
    r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
    r.set_models(s.tris, s.tris_model_count); r.set_scene(s.source_depth, s.source_color, None)
    r.set_observed(obs["aos"], obs["rgb"], None)
    kw = dict(stride=stride, sensor_resolution=0.01, occlusion_threshold=1.0)
    cl = r.render_cuda_multi_unified("CLOUD", poses, pmm, None, **kw)
    for mi in (1, 2, 3, 150):
        g = r.gicp(poses, pmm, None, icp_max_iterations=mi, **kw)
        for p in range(3):
            src = cl["result_cloud"][:, cl["result_cloud_pose_map"] == p].T
            T, it, err = orc.gicp(src, obs["aos"], None, 0, max_iterations=mi)
            Tg = g["transforms"][p].astype(np.float64)
            print(f"maxit {mi} pose {p} n_src {len(src)}: it gpu {g['iterations'][p]} cpu {it}; err gpu {g['error'][p]:.6e} cpu {err:.6e}; "
                  f"dt {np.abs(Tg[9:]-T[9:]).max()*1e3:.4f} mm dr {_angle_deg(Tg[:9].reshape(3,3), T[:9].reshape(3,3)):.5f} deg |t| {np.abs(T[9:]).max()*1e3:.2f} mm")
