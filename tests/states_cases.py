"""Seeded inputs of the caller-glue tests (tests/test_states.py): a 3-DoF table scene and a 6-DoF labelled scene built on
tests/scenes.py, expressed the way sbpl_perception holds them (world = table frame, body-frame camera pose, ObjectModel
metadata, projected / segmented search clouds)."""
from __future__ import annotations

import numpy as np

from perception_b200 import states as S
from scenes import make_case, observed_3dof, observed_6dof

CAM_TO_BODY = np.array([[0, 0, 1, 0], [-1, 0, 0, 0], [0, -1, 0, 0], [0, 0, 0, 1.0]])


def _model_vertices(scene, m):
    lo = int(np.sum(scene.tris_model_count[:m]))
    return scene.tris["v"][lo:lo + int(scene.tris_model_count[m])].reshape(-1, 3)


def make_metas(scene, symmetric=(False, False, False), symmetry_mode=(0, 0, 0), search_resolution=0.0):
    """ObjectModel metadata: the preprocessing transform lifts the model so that its base rests on z = 0 (what
    PreprocessModel does for 3-DoF), min/max are those of the preprocessed mesh (object_model.cpp:236-248, 352-359)."""
    metas = []
    for m in range(len(scene.tris_model_count)):
        v = _model_vertices(scene, m).astype(np.float64)
        pre = np.eye(4, dtype=np.float32)
        pre[2, 3] = np.float32(scene.model_zrest[m])
        vp = (v @ pre[:3, :3].astype(np.float64).T + pre[:3, 3].astype(np.float64)).astype(np.float32)
        metas.append(S.ModelMeta.from_vertices(vp, pre, search_resolution=search_resolution,
                                               symmetric=symmetric[m % len(symmetric)],
                                               symmetry_mode=symmetry_mode[m % len(symmetry_mode)]))
    return metas


def case_3dof(res=0.05, theta_res=np.pi / 4, min_neighbor=20, with_constraint=False, **meta_kw):
    case = make_case()
    s = case.scene
    T_table_cam = np.linalg.inv(s.T_cam_table)  # optical camera frame -> world (table) frame == cam_z_front
    env = S.Env(cam_to_world=T_table_cam @ np.linalg.inv(CAM_TO_BODY), x_min=-0.3, x_max=0.3, y_min=-0.25, y_max=0.25,
                table_height=0.0, res=res, theta_res=theta_res, footprint_tolerance=0.05, use_external_pose_list=0,
                use_cylinder_observed=True, min_neighbor_points_for_valid_pose=min_neighbor,
                min_points_for_constraint_cloud=3)
    obs = observed_3dof(case, 4)
    # observed_cloud_ in the world frame, projected to the table (search_env.cpp:5637-5651)
    world = (obs["aos"].astype(np.float64) @ T_table_cam[:3, :3].T + T_table_cam[:3, 3]).astype(np.float32)
    projected = world.copy()
    projected[:, 2] = np.float32(env.table_height)
    constraint = None
    if with_constraint:
        gt_xy = np.array([(T_table_cam @ T)[:2, 3] for T in s.gt_poses])
        rng = np.random.default_rng(9)
        constraint = np.zeros((24, 3), np.float32)
        constraint[:, :2] = gt_xy[0] + rng.uniform(-0.02, 0.02, size=(24, 2))
    return dict(case=case, env=env, metas=make_metas(s, **meta_kw), obs=obs, projected=projected, constraint=constraint)


def states_3dof(c, n=200, seed=4):
    """Candidate states scattered over (and slightly beyond) the table, plus the ground-truth placements."""
    rng = np.random.default_rng(seed)
    s = c["case"].scene
    n_models = len(c["metas"])
    out = np.zeros(n, S.STATE_DTYPE)
    T_table_cam = np.linalg.inv(s.T_cam_table)
    for i in range(n):
        m = i % n_models
        if i < len(s.gt_poses):
            T = T_table_cam @ s.gt_poses[i]
            x, y, yaw, m = T[0, 3], T[1, 3], np.arctan2(T[1, 0], T[0, 0]), s.gt_models[i]
        else:
            x, y, yaw = rng.uniform(-0.4, 0.4), rng.uniform(-0.35, 0.35), rng.uniform(-7, 7)
        out[i] = S.state_euler(m, x, y, c["env"].table_height, 0.0, 0.0, yaw)
    # a few general Euler triples: the glue must not assume roll = pitch = 0
    for i in range(n - 8, n):
        out[i] = S.state_euler(i % n_models, rng.uniform(-.2, .2), rng.uniform(-.2, .2), 0.01, *rng.uniform(-4, 4, 3))
    return out


def case_6dof(min_neighbor=20):
    case = make_case()
    s = case.scene
    n_labels = len(s.gt_poses)
    env = S.Env(cam_to_world=np.linalg.inv(CAM_TO_BODY), use_external_pose_list=1, res=0.02,
                min_neighbor_points_for_valid_pose=min_neighbor)  # identity cam_z_front: poses live in the camera frame
    obs = observed_6dof(case, 4)
    lab = obs["label"]
    # knn (whole cloud) followed by one segmented cloud per label (search_env.cpp:5593-5605, 5624-5629)
    xyz = [obs["aos"]]
    ranges = [[0, len(obs["aos"])]]
    for l in range(n_labels):
        pts = obs["aos"][lab == l]
        ranges.append([ranges[-1][1], ranges[-1][1] + len(pts)])
        xyz.append(pts)
    metas = []
    for m in range(len(s.tris_model_count)):
        metas.append(S.ModelMeta.from_vertices(_model_vertices(s, m)))
    return dict(case=case, env=env, metas=metas, obs=obs, xyz=np.concatenate(xyz).astype(np.float32),
                ranges=np.array(ranges, np.int32), n_labels=n_labels)


def states_6dof(c, n=120, seed=6):
    """poses.txt-style candidates (x y z qx qy qz qw in the camera frame) around the ground truth."""
    rng = np.random.default_rng(seed)
    s = c["case"].scene
    out = np.zeros(n, S.STATE_DTYPE)
    for i in range(n):
        g = i % len(s.gt_poses)
        T = s.gt_poses[g]
        q = rng.normal(size=4)
        q *= rng.uniform(0.5, 2.0) / np.linalg.norm(q)  # not normalised on purpose: GetTransform normalises
        t = T[:3, 3] + rng.normal(0, 0.03 if i % 5 else 0.3, 3)
        out[i] = S.state_quat(s.gt_models[g], t[0], t[1], t[2], q[1], q[2], q[3], q[0], segmentation_label_id=g + 1)
    return out
