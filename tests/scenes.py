"""Seeded test scenes shared by the CPU and GPU tests.  Scene depth maps are rasterised with the oracle (test infra)."""
from __future__ import annotations

from dataclasses import dataclass
from functools import lru_cache

import numpy as np

from oracle_lib import Oracle
from perception_b200 import synth
from perception_b200.renderer import pose_to_mat4x4


@dataclass
class Case:
    scene: synth.Scene
    proj: np.ndarray
    width: int
    height: int
    cx: float
    cy: float
    fx: float
    fy: float


def oracle_render_fn(width, height, proj):
    orc = Oracle()

    def fn(tris, counts, poses, pmm):
        d, c = orc.render(tris, counts, poses, pmm, width, height, proj)
        return d, c

    return fn


@lru_cache(maxsize=8)
def make_case(n_tris: int = 600, n_instances: int = 3, seed: int = 42, width: int = 640, height: int = 480,
              with_box: bool = True) -> Case:
    """Camera scaled from the reference's 640x480 intrinsics when a smaller image is asked for."""
    sx, sy = width / synth.WIDTH, height / synth.HEIGHT
    fx, fy, cx, cy = synth.FX * sx, synth.FY * sy, synth.CX * sx, synth.CY * sy
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    proj = Oracle().compute_proj(K, width, height)
    models = [synth.make_mesh(n_tris, seed=1000, r0=0.08), synth.make_mesh(max(n_tris // 2, 24), seed=1001, r0=0.06)]
    if with_box:
        models.append(synth.make_box_mesh((0.05, 0.035, 0.06)))
    scene = synth.make_scene(models, oracle_render_fn(width, height, proj), n_instances=n_instances, seed=seed,
                             width=width, height=height)
    return Case(scene, proj, width, height, float(np.float32(cx)), float(np.float32(cy)), float(np.float32(fx)),
                float(np.float32(fy)))


def observed_3dof(case: Case, stride: int, orc=None):
    """Observed cloud the way SetInput builds it for 3-DoF (search_env.cpp:5993-6017): table-frame bounds filter."""
    orc = orc or Oracle()
    s = case.scene
    T_tc = np.linalg.inv(s.T_cam_table)
    return orc.depth2cloud(s.source_depth, s.source_color, case.width, case.height, 1, case.cx, case.cy, case.fx, case.fy,
                           synth.DEPTH_FACTOR, stride, bounds=synth.table_bounds(s), cam_tf=T_tc)


def observed_6dof(case: Case, stride: int, orc=None):
    """Observed cloud for 6-DoF: label-mask filter, labels = mask-1."""
    orc = orc or Oracle()
    s = case.scene
    return orc.depth2cloud(s.source_depth, s.source_color, case.width, case.height, 1, case.cx, case.cy, case.fx, case.fy,
                           synth.DEPTH_FACTOR, stride, label_mask=s.source_label)


def poses_near_gt(case: Case, n: int, seed: int = 5, trans_sigma: float = 0.01, rot_sigma: float = 0.15):
    """n candidate poses: ground-truth instances with seeded perturbations (+ a few far-off ones).  Returns
    (poses (n,4,4) float32 x100, pose_model_map (n,), pose_label (n,) = instance index)."""
    rng = np.random.default_rng(seed)
    s = case.scene
    poses, pmm, lab = [], [], []
    for i in range(n):
        g = i % len(s.gt_poses)
        T = s.gt_poses[g].copy()
        if i >= len(s.gt_poses):  # first ones are exact ground truth
            w = rng.normal(0, rot_sigma, 3)
            th = np.linalg.norm(w)
            k = w / th if th > 0 else np.array([0, 0, 1.0])
            Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
            R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx
            T[:3, :3] = T[:3, :3] @ R
            scale = 10.0 if i % 11 == 10 else 1.0
            T[:3, 3] += rng.normal(0, trans_sigma * scale, 3)
        poses.append(pose_to_mat4x4(T))
        pmm.append(s.gt_models[g])
        lab.append(g)
    return np.stack(poses), np.array(pmm, np.int32), np.array(lab, np.int32)
