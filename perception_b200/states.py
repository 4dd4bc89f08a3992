"""Host-side types of the caller glue (include/perch_b200.h, "caller glue" section).

Mirrors what sbpl_perception keeps around render_cuda_multi_unified: ``ContPose`` / ``ObjectState``
(object_state.h, object_state.cpp:17-97), the ``ObjectModel`` geometry the glue reads (object_model.cpp:326-466) and the
``EnvParams`` / ``PerchParams`` fields of search_env.cpp:140-216.  Plain numpy; the arithmetic that matters for parity
(pose matrices, radius counts, validity) runs on the GPU through ``Renderer`` -- nothing here is a CPU path for it.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field
from typing import Sequence

import numpy as np

from ._capi import perch_env, perch_model_meta

# perch_state: 10 doubles + 2 int32 = 88 bytes
STATE_DTYPE = np.dtype(
    [(n, "<f8") for n in ("x", "y", "z", "roll", "pitch", "yaw", "qx", "qy", "qz", "qw")]
    + [("model_id", "<i4"), ("segmentation_label_id", "<i4")]
)
assert STATE_DTYPE.itemsize == 88

K_MESH_ADDITIVE_INFLATION = 0.01  # object_model.cpp:43


def normalize_angle_positive(a: float) -> float:
    """angles::normalize_angle_positive, as ContPose's Euler constructor applies it (object_state.cpp:17-22)."""
    return math.fmod(math.fmod(a, 2.0 * math.pi) + 2.0 * math.pi, 2.0 * math.pi)


def state_euler(model_id: int, x, y, z, roll, pitch, yaw, segmentation_label_id: int = -1) -> np.ndarray:
    """ObjectState(id, symmetric, ContPose(x, y, z, roll, pitch, yaw)[, segmentation_label_id])."""
    s = np.zeros((), STATE_DTYPE)
    s["x"], s["y"], s["z"] = x, y, z
    s["roll"], s["pitch"], s["yaw"] = (normalize_angle_positive(v) for v in (roll, pitch, yaw))
    s["model_id"], s["segmentation_label_id"] = model_id, segmentation_label_id
    return s


def state_quat(model_id: int, x, y, z, qx, qy, qz, qw, segmentation_label_id: int = -1) -> np.ndarray:
    """ObjectState(id, symmetric, ContPose(x, y, z, qx, qy, qz, qw), segmentation_label_id): the poses.txt line of the
    6-DoF flow (search_env.cpp:7116-7146).  The Euler fields the constructor also fills (object_state.cpp:27-40) are
    not read by ContPose::GetTransform when the quaternion is non-zero and stay 0 here."""
    s = np.zeros((), STATE_DTYPE)
    s["x"], s["y"], s["z"] = x, y, z
    s["qx"], s["qy"], s["qz"], s["qw"] = qx, qy, qz, qw
    s["model_id"], s["segmentation_label_id"] = model_id, segmentation_label_id
    return s


def states_from_pose_lines(model_id: int, lines: np.ndarray, segmentation_label_id: int) -> np.ndarray:
    """poses.txt rows `x y z qx qy qz qw` (search_env.cpp:7116-7128) -> states of one model."""
    a = np.asarray(lines, np.float64).reshape(-1, 7)
    s = np.zeros(len(a), STATE_DTYPE)
    for k, n in enumerate(("x", "y", "z", "qx", "qy", "qz", "qw")):
        s[n] = a[:, k]
    s["model_id"] = model_id
    s["segmentation_label_id"] = segmentation_label_id
    return s


def load_pose_list(path_or_text, model_id: int, segmentation_label_id: int) -> np.ndarray:
    """perch_parse_pose_list: a poses.txt file (or its text) of the 6-DoF flow (search_env.cpp:7099-7146) -> states."""
    import ctypes as C
    import os
    from . import _capi
    lib = _capi.load_library()
    if isinstance(path_or_text, (bytes, bytearray)):
        text = bytes(path_or_text)
    elif isinstance(path_or_text, str) and "\n" not in path_or_text and os.path.exists(path_or_text):
        text = open(path_or_text, "rb").read()
    else:
        text = str(path_or_text).encode()
    n = C.c_int(0)
    _capi.check(lib, lib.perch_parse_pose_list(text, len(text), model_id, segmentation_label_id, None, 0, C.byref(n)))
    out = np.zeros(n.value, STATE_DTYPE)
    _capi.check(lib, lib.perch_parse_pose_list(text, len(text), model_id, segmentation_label_id, out.ctypes.data_as(C.c_void_p),
                                               n.value, C.byref(n)))
    return out


def convex_hull_xy(points_xy: np.ndarray) -> np.ndarray:
    """Convex hull of the projected mesh vertices (object_model.cpp:361-372 uses pcl::ConvexHull, dimension 2):
    Andrew's monotone chain, counter-clockwise, no repeated end point (an implicitly closed polygon)."""
    pts = np.unique(np.asarray(points_xy, np.float64).reshape(-1, 2), axis=0)
    if len(pts) <= 2:
        return pts

    def half(seq):
        h = []
        for p in seq:
            while len(h) >= 2 and ((h[-1][0] - h[-2][0]) * (p[1] - h[-2][1]) - (h[-1][1] - h[-2][1]) * (p[0] - h[-2][0])) <= 0:
                h.pop()
            h.append(p)
        return h

    lower, upper = half(pts), half(pts[::-1])
    return np.array(lower[:-1] + upper[:-1])


@dataclass
class ModelMeta:
    """What the glue reads from ObjectModel / ModelMetaData."""

    preprocessing_transform: np.ndarray = field(default_factory=lambda: np.eye(4, dtype=np.float32))
    circumscribed_radius: float = 0.0
    circumscribed_radius_3d: float = 0.0
    inscribed_radius: float = 0.0
    inflation_factor: float = 1.0
    search_resolution: float = 0.0
    symmetric: bool = False
    symmetry_mode: int = 0
    footprint: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))

    @classmethod
    def from_vertices(cls, vertices: np.ndarray, preprocessing_transform=None, search_resolution=0.0, symmetric=False,
                      symmetry_mode=0) -> "ModelMeta":
        """ObjectModel::SetObjectProperties (object_model.cpp:326-384) for a (preprocessed) vertex cloud: float
        min/max (pcl::getMinMax3D), the radii of 456-466 in double, inflation factor 1 + 0.01 / inscribed radius."""
        v = np.asarray(vertices, np.float32).reshape(-1, 3)
        mn, mx = v.min(axis=0).astype(np.float64), v.max(axis=0).astype(np.float64)
        ex = np.abs(mx - mn)
        insc = min(ex[0], ex[1]) / 2.0
        hull = convex_hull_xy(v[:, :2])
        fp = np.zeros((len(hull), 3), np.float32)
        fp[:, :2] = hull
        return cls(
            preprocessing_transform=np.eye(4, dtype=np.float32) if preprocessing_transform is None
            else np.asarray(preprocessing_transform, np.float32).reshape(4, 4),
            circumscribed_radius=max(ex[0], ex[1]) / 2.0,
            circumscribed_radius_3d=max(ex[0], max(ex[1], ex[2])) / 2.0,
            inscribed_radius=insc,
            inflation_factor=1.0 if insc < 1e-5 else 1.0 + (K_MESH_ADDITIVE_INFLATION / insc),
            search_resolution=search_resolution, symmetric=symmetric, symmetry_mode=symmetry_mode, footprint=fp)


@dataclass
class Env:
    """EnvParams / PerchParams fields the glue reads (search_env.cpp:140-216); defaults are the reference's."""

    cam_to_world: np.ndarray = field(default_factory=lambda: np.eye(4))
    x_min: float = 0.0
    x_max: float = 0.0
    y_min: float = 0.0
    y_max: float = 0.0
    table_height: float = 0.0
    res: float = 0.1
    theta_res: float = math.pi / 8
    footprint_tolerance: float = 0.05
    use_external_pose_list: int = 0
    use_cylinder_observed: bool = True
    min_neighbor_points_for_valid_pose: int = 50
    use_adaptive_resolution: bool = False
    use_model_specific_search_resolution: bool = False
    min_points_for_constraint_cloud: int = 50


def env_struct(env: Env) -> perch_env:
    e = perch_env()
    e.cam_to_world = (C.c_double * 16)(*np.asarray(env.cam_to_world, np.float64).reshape(16))
    for n in ("x_min", "x_max", "y_min", "y_max", "table_height", "res", "theta_res", "footprint_tolerance"):
        setattr(e, n, float(getattr(env, n)))
    for n in ("use_external_pose_list", "use_cylinder_observed", "min_neighbor_points_for_valid_pose",
              "use_adaptive_resolution", "use_model_specific_search_resolution", "min_points_for_constraint_cloud"):
        setattr(e, n, int(getattr(env, n)))
    return e


def meta_array(metas: Sequence[ModelMeta], footprints=None):
    """-> (ctypes array of perch_model_meta, (N,3) float32 footprint points of all models back to back)."""
    arr = (perch_model_meta * len(metas))()
    fps, off = [], 0
    for i, m in enumerate(metas):
        arr[i].preprocessing_transform = (C.c_float * 16)(*np.asarray(m.preprocessing_transform, np.float32).reshape(16))
        arr[i].circumscribed_radius = float(m.circumscribed_radius)
        arr[i].circumscribed_radius_3d = float(m.circumscribed_radius_3d)
        arr[i].inscribed_radius = float(m.inscribed_radius)
        arr[i].inflation_factor = float(m.inflation_factor)
        arr[i].search_resolution = float(m.search_resolution)
        arr[i].symmetric = int(bool(m.symmetric))
        arr[i].symmetry_mode = int(m.symmetry_mode)
        fp = np.asarray(m.footprint if footprints is None else footprints[i], np.float32).reshape(-1, 3)
        arr[i].footprint_offset = off
        arr[i].footprint_count = len(fp)
        off += len(fp)
        fps.append(fp)
    all_fp = np.ascontiguousarray(np.concatenate(fps) if fps else np.zeros((0, 3), np.float32), dtype=np.float32)
    return arr, all_fp
