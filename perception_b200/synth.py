"""Deterministic synthetic inputs of the shapes BASELINE.json names: meshes, a tabletop RGB-D
scene, the observed cloud inputs and candidate pose sets.  numpy only -- this is input generation, not the hot path.
The scene's depth image is produced by whichever rasterizer the caller passes in (`render_fn`): the product
renderer on a GPU box, the oracle in CPU-only tests.

Conventions (reference): camera frame x right / y down / z forward, metres; poses are 4x4 object->camera, converted
with ``pose_to_mat4x4(T, 100)`` so rendered depth is integer centimetres; depth_factor = 100.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Tuple

import numpy as np

from .renderer import TRIANGLE_DTYPE, make_triangles, pose_to_mat4x4

# object_recognizer.cpp:64-75 (kinect intrinsics used throughout the reference's configs)
WIDTH, HEIGHT = 640, 480
FX = FY = 576.09757860
CX, CY = 321.06398107, 242.97676897
DEPTH_FACTOR = 100.0


def camera_K() -> np.ndarray:
    return np.array([[FX, 0, CX], [0, FY, CY], [0, 0, 1]], np.float32)


# ---- meshes ------------------------------------------------------------------------------------------------------
def make_mesh(n_tris: int, seed: int, r0: Optional[float] = None, shuffle_seed: int = 7) -> np.ndarray:
    """Closed genus-0 surface r(theta,phi) = r0*(1 + 0.25*sum a_k Y_k) as a UV sphere with exactly n_tris triangles
    (n_tris = 2*n_lon*(n_lat-1); 5 000 -> 50x51, 20 000 -> 100x101), per-triangle colour from seeded smooth noise,
    triangle order shuffled.  Returns reference-layout Triangles (metres)."""
    rng = np.random.default_rng(seed)
    # choose n_lon, n_lat with 2*n_lon*(n_lat-1) == n_tris, as square as possible
    best = None
    for n_lon in range(3, n_tris):
        if n_tris % (2 * n_lon):
            continue
        n_lat = n_tris // (2 * n_lon) + 1
        if n_lat < 3:
            break
        score = abs(n_lon - n_lat)
        if best is None or score < best[0]:
            best = (score, n_lon, n_lat)
    if best is None:
        raise ValueError(f"cannot tessellate a UV sphere into exactly {n_tris} triangles")
    _, n_lon, n_lat = best
    if r0 is None:
        r0 = float(rng.uniform(0.04, 0.12))
    amp = rng.uniform(-1, 1, size=(6,)) / 6.0
    ph = rng.uniform(0, 2 * np.pi, size=(6,))

    def radius(theta, phi):
        s = (amp[0] * np.cos(theta + ph[0]) + amp[1] * np.sin(theta) * np.cos(phi + ph[1]) +
             amp[2] * np.sin(theta) ** 2 * np.cos(2 * phi + ph[2]) + amp[3] * np.cos(2 * theta + ph[3]) +
             amp[4] * np.sin(2 * theta) * np.sin(phi + ph[4]) + amp[5] * np.sin(theta) ** 3 * np.cos(3 * phi + ph[5]))
        return r0 * (1 + 0.25 * 1.5 * s)

    # vertices: north pole, (n_lat-1) rings... rings at theta_i = pi*i/n_lat for i=1..n_lat-1 -> n_lat-1 rings
    thetas = np.pi * np.arange(1, n_lat) / n_lat
    phis = 2 * np.pi * np.arange(n_lon) / n_lon
    T, Ph = np.meshgrid(thetas, phis, indexing="ij")
    R = radius(T, Ph)
    ring = np.stack([R * np.sin(T) * np.cos(Ph), R * np.sin(T) * np.sin(Ph), R * np.cos(T)], -1)  # (n_lat-1, n_lon, 3)
    north = np.array([0, 0, radius(0.0, 0.0)])
    south = np.array([0, 0, -radius(np.pi, 0.0)])
    tris = []
    j = np.arange(n_lon)
    jn = (j + 1) % n_lon
    # caps: n_lon triangles each
    tris.append(np.stack([np.broadcast_to(north, (n_lon, 3)), ring[0, j], ring[0, jn]], 1))
    tris.append(np.stack([np.broadcast_to(south, (n_lon, 3)), ring[-1, jn], ring[-1, j]], 1))
    # bands: (n_lat-2) bands x 2 triangles x n_lon
    for i in range(n_lat - 2):
        a, b, c, d = ring[i, j], ring[i, jn], ring[i + 1, j], ring[i + 1, jn]
        tris.append(np.stack([a, c, b], 1))
        tris.append(np.stack([b, c, d], 1))
    V = np.concatenate(tris, 0).astype(np.float32)
    assert len(V) == n_tris, (len(V), n_tris)
    cen = V.mean(1)
    f = rng.uniform(20, 60, size=(3, 3))
    p0 = rng.uniform(0, 2 * np.pi, size=(3,))
    col = 128 + 100 * np.sin(cen @ f.T + p0)
    col = np.clip(col, 0, 255).astype(np.uint64)
    order = np.random.default_rng(shuffle_seed).permutation(n_tris)
    return make_triangles(V[order], col[order])


def make_box_mesh(half: Tuple[float, float, float], color=(200, 120, 40)) -> np.ndarray:
    """12-triangle box: the low-poly / huge-triangle edge case."""
    hx, hy, hz = half
    c = np.array([[sx * hx, sy * hy, sz * hz] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)], np.float32)
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    V = []
    for q in quads:
        V.append([c[q[0]], c[q[1]], c[q[2]]])
        V.append([c[q[0]], c[q[2]], c[q[3]]])
    V = np.array(V, np.float32)
    col = np.tile(np.array(color, np.uint64), (len(V), 1))
    col[:, 0] = (col[:, 0] + 10 * np.arange(len(V))) % 256
    return make_triangles(V, col)


def concat_models(models: List[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """-> (tris, tris_model_count) as EnvObjectRecognition builds them (search_env.cpp:299-305)."""
    return np.concatenate(models).astype(TRIANGLE_DTYPE), np.array([len(m) for m in models], np.int32)


# ---- frames / poses -----------------------------------------------------------------------------------------------
def table_frame(distance: float = 0.9, tilt_deg: float = 35.0) -> np.ndarray:
    """T_cam_table: table origin `distance` along the optical axis, camera looking `tilt_deg` down onto the plane."""
    c, s = np.cos(np.radians(tilt_deg)), np.sin(np.radians(tilt_deg))
    T = np.eye(4)
    T[:3, 0] = (1, 0, 0)
    T[:3, 1] = (0, -s, c)
    T[:3, 2] = (0, -c, -s)  # table normal, towards the camera
    T[:3, 3] = (0, 0, distance)
    return T


def rot_z(a: float) -> np.ndarray:
    c, s = np.cos(a), np.sin(a)
    R = np.eye(4)
    R[:2, :2] = [[c, -s], [s, c]]
    return R


def pose_on_table(T_cam_table: np.ndarray, x: float, y: float, yaw: float, z_rest: float) -> np.ndarray:
    T = rot_z(yaw)
    T[:3, 3] = (x, y, z_rest)
    return T_cam_table @ T


def random_rotation(rng) -> np.ndarray:
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def fibonacci_views(n: int) -> np.ndarray:
    """n unit view directions (fat_pose_image.py:903-966 samples viewpoints on a Fibonacci sphere)."""
    i = np.arange(n) + 0.5
    phi = np.arccos(1 - 2 * i / n)
    th = np.pi * (1 + 5 ** 0.5) * i
    return np.stack([np.cos(th) * np.sin(phi), np.sin(th) * np.sin(phi), np.cos(phi)], -1)


def rotation_from_view(v: np.ndarray, inplane: float) -> np.ndarray:
    z = v / np.linalg.norm(v)
    a = np.array([0, 0, 1.0]) if abs(z[2]) < 0.9 else np.array([1.0, 0, 0])
    x = np.cross(a, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z], 1)
    return R @ rot_z(inplane)[:3, :3]


# ---- scene -----------------------------------------------------------------------------------------------------------
@dataclass
class Scene:
    width: int
    height: int
    K: np.ndarray
    T_cam_table: np.ndarray
    tris: np.ndarray
    tris_model_count: np.ndarray
    gt_poses: List[np.ndarray]            # object->camera, metres
    gt_models: List[int]
    source_depth: np.ndarray              # (H,W) int32 cm
    source_color: np.ndarray              # (3,H,W) uint8
    source_label: np.ndarray              # (H,W) uint8, 0 = background/table, i+1 = gt instance i
    model_radius: np.ndarray              # circumscribed xy radius per model (metres)
    model_zrest: np.ndarray               # -min z per model
    extra: dict = field(default_factory=dict)


RenderFn = Callable[[np.ndarray, np.ndarray, np.ndarray, np.ndarray], Tuple[np.ndarray, np.ndarray]]


def make_scene(models: List[np.ndarray], render_fn: RenderFn, n_instances: int = 4, seed: int = 42,
               width: int = WIDTH, height: int = HEIGHT, noise_cm: float = 0.3, dropout: float = 0.02,
               table_half: float = 0.45) -> Scene:
    """Table plane + n_instances objects (model i % len(models)), rendered with render_fn into one int32 cm depth map +
    RGB + label mask; additive depth noise round(N(0, noise_cm)) and `dropout` zeroed pixels."""
    rng = np.random.default_rng(seed)
    T_ct = table_frame()
    radius = np.array([float(np.linalg.norm(m["v"][..., :2], axis=-1).max()) for m in models])
    zrest = np.array([float(-m["v"][..., 2].min()) for m in models])
    table = make_triangles(
        np.array([[[-1, -1, 0], [1, -1, 0], [1, 1, 0]], [[-1, -1, 0], [1, 1, 0], [-1, 1, 0]]], np.float32) * 0.6,
        np.array([[90, 90, 90], [90, 90, 90]]))
    all_models = models + [table]
    tris, counts = concat_models(all_models)
    gt_poses, gt_models, placed = [], [], []
    tries = 0
    while len(gt_poses) < n_instances and tries < 1000:
        tries += 1
        m = len(gt_poses) % len(models)
        x, y = rng.uniform(-0.25, 0.25), rng.uniform(-0.15, 0.2)
        if any(np.hypot(x - px, y - py) < radius[m] + pr + 0.01 for px, py, pr in placed):
            continue
        placed.append((x, y, radius[m]))
        gt_poses.append(pose_on_table(T_ct, x, y, rng.uniform(0, 2 * np.pi), zrest[m]))
        gt_models.append(m)
    poses = [pose_to_mat4x4(T) for T in gt_poses] + [pose_to_mat4x4(T_ct)]
    pmm = np.array(gt_models + [len(models)], np.int32)
    depth, color = render_fn(tris, counts, np.stack(poses), pmm)  # (P,H,W), (3,P,H,W)
    big = np.where(depth > 0, depth, np.iinfo(np.int32).max)
    win = big.argmin(0)
    d = np.take_along_axis(big, win[None], 0)[0]
    valid = d < np.iinfo(np.int32).max
    src_depth = np.where(valid, d, 0).astype(np.int32)
    src_color = np.stack([np.where(valid, np.take_along_axis(color[k], win[None], 0)[0], 0) for k in range(3)]).astype(np.uint8)
    label = np.where(valid & (win < len(gt_poses)), win + 1, 0).astype(np.uint8)
    noise = np.rint(rng.normal(0, noise_cm, size=src_depth.shape)).astype(np.int32)
    src_depth = np.where(src_depth > 0, np.maximum(src_depth + noise, 1), 0).astype(np.int32)
    drop = rng.uniform(size=src_depth.shape) < dropout
    src_depth[drop] = 0
    label[drop] = 0
    mtris, mcounts = concat_models(models)
    return Scene(width, height, camera_K(), T_ct, mtris, mcounts, gt_poses, gt_models, src_depth, src_color, label,
                 radius, zrest, {"table_half": table_half})


def table_bounds(scene: Scene, zmin: float = 0.005, zmax: float = 0.5) -> np.ndarray:
    """{xmax,xmin,ymax,ymin,zmax,zmin} in the table frame (compute_point_clouds.cuh:86-88); zmin > 0 drops the table."""
    h = scene.extra.get("table_half", 0.45)
    return np.array([h, -h, h, -h, zmax, zmin], np.float64)


# ---- candidate poses ---------------------------------------------------------------------------------------------------
def poses_3dof(scene: Scene, model: int, n: int, step: float = 0.02, n_yaw: int = 16, seed: int = 1) -> np.ndarray:
    """Grid x,y over the table bounds x n_yaw yaws (search_env.cpp:7276-7311), keeping the candidates nearest to any
    ground-truth instance first so the batch contains good and bad hypotheses; returns (n,4,4) mat4x4 (x100)."""
    h = scene.extra.get("table_half", 0.45)
    xs = np.arange(-h, h + 1e-9, step)
    g = np.stack(np.meshgrid(xs, xs, indexing="ij"), -1).reshape(-1, 2)
    gt_xy = np.array([(np.linalg.inv(scene.T_cam_table) @ T)[:2, 3] for T in scene.gt_poses])
    dist = np.min(np.linalg.norm(g[:, None] - gt_xy[None], axis=-1), 1)
    order = np.argsort(dist, kind="stable")
    out = []
    for gi in order:
        for k in range(n_yaw):
            out.append(pose_to_mat4x4(pose_on_table(scene.T_cam_table, g[gi, 0], g[gi, 1], 2 * np.pi * k / n_yaw,
                                                    scene.model_zrest[model])))
            if len(out) == n:
                return np.stack(out)
    while len(out) < n:
        out.append(out[len(out) % max(1, len(out))])
    return np.stack(out)


def poses_6dof(center_cam: np.ndarray, n: int, n_views: int = 512, n_inplane: int = 32, jitter: float = 0.01,
               seed: int = 3) -> np.ndarray:
    """n_views Fibonacci-sphere views x n_inplane in-plane rotations, translation = center +- U(-jitter,jitter) metres;
    returns (n,4,4) mat4x4 (x100)."""
    rng = np.random.default_rng(seed)
    views = fibonacci_views(n_views)
    out = np.zeros((n, 4, 4), np.float32)
    for i in range(n):
        v, k = (i // n_inplane) % n_views, i % n_inplane
        T = np.eye(4)
        T[:3, :3] = rotation_from_view(views[v], 2 * np.pi * k / n_inplane)
        T[:3, 3] = center_cam + rng.uniform(-jitter, jitter, size=3)
        out[i] = pose_to_mat4x4(T)
    return out


def cylinder_counts(scene: Scene, obs_xyz: np.ndarray, poses: np.ndarray, pose_model_map: np.ndarray) -> np.ndarray:
    """pose_observed_points_total for 3-DoF (search_env.cpp:1593-1614): observed points whose table-frame (x,y) lies
    within the model's circumscribed radius of the pose's (x,y).  Caller-side glue, numpy."""
    T_tc = np.linalg.inv(scene.T_cam_table)
    pts = obs_xyz.astype(np.float64) @ T_tc[:3, :3].T + T_tc[:3, 3]
    out = np.zeros(len(poses), np.float32)
    xy = pts[:, :2]
    if len(poses) > 4096:
        # large synthetic batches (bench workloads): the same count through a k-d tree
        from scipy.spatial import cKDTree
        t = np.asarray(poses, np.float64).reshape(-1, 4, 4)[:, :3, 3] / 100.0
        c = (t @ T_tc[:3, :3].T + T_tc[:3, 3])[:, :2]
        r = np.asarray(scene.model_radius, np.float64)[np.asarray(pose_model_map)]
        return np.asarray(cKDTree(xy).query_ball_point(c, r, return_length=True), np.float32)
    for i, (P, m) in enumerate(zip(poses, pose_model_map)):
        t = np.asarray(P, np.float64)[:3, 3] / 100.0
        c = (T_tc[:3, :3] @ t + T_tc[:3, 3])[:2]
        out[i] = np.count_nonzero(np.sum((xy - c) ** 2, axis=1) <= scene.model_radius[m] ** 2)
    return out
