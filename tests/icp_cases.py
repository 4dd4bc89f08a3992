"""Seeded inputs of the cuda_icp parity cases: a 160x120 synthetic scene (int32 cm depth) and model clouds = scene
points of one object moved by a small rigid transform (what ICP has to undo)."""
import numpy as np

from perception_b200 import synth
from scenes import make_case

W, H = 640, 480


def scene():
    case = make_case(600, 3, 42, W, H)
    K = np.array([[case.fx, 0, case.cx], [0, case.fy, case.cy], [0, 0, 1]], np.float32)
    return case, K


def _rot(axis, ang):
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    Kx = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * Kx + (1 - np.cos(ang)) * Kx @ Kx


def clouds(scene_pcd, label_img, n_clouds=4, seed=0):
    """Clouds sampled from the scene points of labelled objects, displaced by up to 4 mm / 1.5 deg about their centroid."""
    rng = np.random.default_rng(seed)
    out, moves = [], []
    labels = [l for l in np.unique(label_img) if l > 0]
    for i in range(n_clouds):
        l = labels[i % len(labels)]
        pts = scene_pcd[(label_img.reshape(-1) == l) & (scene_pcd[:, 2] > 0)].astype(np.float64)
        pts = pts[rng.choice(len(pts), min(len(pts), 400), replace=False)]
        R = _rot(rng.normal(size=3), np.deg2rad(rng.uniform(0.3, 1.5)))
        c = pts.mean(0)
        t = rng.uniform(-0.004, 0.004, 3)
        moved = (pts - c) @ R.T + c + t
        out.append(moved.astype(np.float32))
        moves.append((R, c - R @ c + t))
    return out, moves
