"""Host mirror of the reference's ``cuda_renderer`` interface on top of libperch_b200.so.

Names and argument meaning follow cuda_renderer/include/cuda_renderer/renderer.h:
``compute_proj`` (:87), ``depth2cloud_global`` (:131-153), ``render_cuda_multi_unified`` (:221-268).
``Renderer`` keeps models / camera / scene / observed cloud resident on the GPU between calls (the reference
re-uploads them on every call, renderer.cu:1532-1544); the module-level functions with the reference's full argument
lists are stateless wrappers over it.  Arrays are numpy; nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _capi
from ._capi import PerchError, perch_outputs, perch_params  # noqa: F401

# cuda_renderer::Model::Triangle (model.h:62-73): 3 x float3, 4 bytes padding, 3 x size_t colour == 64 bytes
TRIANGLE_DTYPE = np.dtype(
    {"names": ["v", "color"], "formats": [("<f4", (3, 3)), ("<u8", (3,))], "offsets": [0, 40], "itemsize": 64}
)


def make_triangles(vertices: np.ndarray, colors: np.ndarray) -> np.ndarray:
    """(T,3,3) float32 vertices + (T,3) colours -> array of 64-byte reference Triangles."""
    t = np.zeros(len(vertices), dtype=TRIANGLE_DTYPE)
    t["v"] = np.asarray(vertices, dtype=np.float32)
    t["color"] = np.asarray(colors, dtype=np.uint64)
    return t


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dtype) -> Optional[np.ndarray]:
    if a is None:
        return None
    return np.ascontiguousarray(a, dtype=dtype)


def compute_proj(K: np.ndarray, width: int, height: int, near: float = 10.0, far: float = 10000.0) -> np.ndarray:
    """cuda_renderer::compute_proj (renderer.h:87): pinhole K (3x3) -> 4x4 projection (row-major mat4x4)."""
    lib = _capi.load_library()
    Kc = _c(K, np.float32).reshape(9)
    out = np.zeros(16, dtype=np.float32)
    _capi.check(lib, lib.perch_compute_proj(_ptr(Kc), width, height, near, far, _ptr(out)))
    return out.reshape(4, 4)


def pose_to_mat4x4(pose_in_cam: np.ndarray, scale_factor: float = 100.0) -> np.ndarray:
    """Model::mat4x4::init_from_eigen (model.h:89-107): double 4x4 -> float, top three rows scaled (m -> cm)."""
    m = np.asarray(pose_in_cam, dtype=np.float64).copy()
    m[..., :3, :] *= scale_factor
    return m.astype(np.float32)


class Renderer:
    """One GPU context (perch_ctx).  Not thread-safe; one per device."""

    def __init__(self, device: int = 0):
        self._lib = _capi.load_library()
        h = C.c_void_p()
        _capi.check(self._lib, self._lib.perch_create(device, C.byref(h)))
        self._h = h
        self.width = self.height = 0
        self.n_models = 0

    # -- lifetime ----------------------------------------------------------------------------------------------
    _n_generated = 0

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.perch_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, code: int) -> None:
        _capi.check(self._lib, code)

    # -- state ---------------------------------------------------------------------------------------------------
    def set_camera(self, width, height, proj_mat, cx, cy, fx, fy, depth_factor=100.0) -> None:
        p = _c(proj_mat, np.float32).reshape(16)
        self._ck(self._lib.perch_set_camera(self._h, width, height, _ptr(p), cx, cy, fx, fy, depth_factor))
        self.width, self.height = int(width), int(height)

    def set_models(self, tris: np.ndarray, tris_model_count: Sequence[int]) -> None:
        """@tris / @tris_model_count of render_cuda_multi_unified (renderer.cu:1433,1436)."""
        t = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        self._ck(self._lib.perch_set_models(self._h, _ptr(t), len(t), _ptr(cnt), len(cnt)))
        self.n_models = len(cnt)

    def set_scene(self, source_depth, source_color=None, source_mask_label=None) -> None:
        """@source_depth (int32 H*W, cm), @source_color (3 planes), @source_mask_label (renderer.cu:1438-1442)."""
        d = _c(source_depth, np.int32).reshape(-1)
        assert d.size == self.width * self.height, "source_depth size != width*height"
        planes = [None, None, None]
        if source_color is not None:
            planes = [_c(source_color[k], np.uint8).reshape(-1) for k in range(3)]
        lab = None if source_mask_label is None or len(source_mask_label) == 0 else _c(source_mask_label, np.uint8).reshape(-1)
        self._ck(self._lib.perch_set_scene(self._h, _ptr(d), _ptr(planes[0]), _ptr(planes[1]), _ptr(planes[2]), _ptr(lab)))

    def set_observed(self, observed_xyz, observed_color=None, observed_label=None) -> None:
        """@observed_depth_eigen (N,3), @observed_color (3,N), @result_observed_cloud_label (renderer.cu:1451-1456)."""
        xyz = _c(observed_xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        col = None if observed_color is None else _c(observed_color, np.uint8).reshape(3, n)
        lab = None if observed_label is None else _c(observed_label, np.int32).reshape(n)
        self._ck(self._lib.perch_set_observed(self._h, _ptr(xyz), _ptr(col), _ptr(lab), n))
        self.n_obs = n

    # -- depth image -> cloud ---------------------------------------------------------------------------------------
    def depth2cloud_global(self, depth_data, color_data, width, height, num_poses, cx, cy, fx, fy, depth_factor, stride,
                           label_mask_data=None, observed_cloud_bounds=None, camera_transform=None, want_dc_index=False):
        """cuda_renderer::depth2cloud_global (renderer.h:131-153).  Returns a dict with the reference's outputs:
        result_cloud_eigen (N,3), result_cloud (3,N), result_cloud_color (3,N), dc_index, rendered_cloud_point_num,
        cloud_pose_map, result_observed_cloud_label (labels - 1)."""
        d = _c(depth_data, np.int32).reshape(-1)
        planes = [None, None, None]
        if color_data is not None:
            planes = [_c(color_data[k], np.uint8).reshape(-1) for k in range(3)]
        lab = None if label_mask_data is None or len(label_mask_data) == 0 else _c(label_mask_data, np.uint8).reshape(-1)
        bnd = None if observed_cloud_bounds is None or len(observed_cloud_bounds) == 0 else _c(observed_cloud_bounds, np.float64)
        # Eigen::Matrix4f is column-major
        tf = None if camera_transform is None else np.ascontiguousarray(np.asarray(camera_transform, np.float32).T).reshape(16)
        n = C.c_int(0)
        args = (width, height, num_poses, cx, cy, fx, fy, depth_factor, stride, _ptr(lab), _ptr(bnd), _ptr(tf))
        self._ck(self._lib.perch_depth2cloud(self._h, _ptr(d), _ptr(planes[0]), _ptr(planes[1]), _ptr(planes[2]), *args,
                                             C.byref(n), None, None, None, None, None, None))
        N = n.value
        soa = np.zeros((3, N), np.float32)
        aos = np.zeros((N, 3), np.float32)
        col = np.zeros((3, N), np.uint8)
        pmap = np.zeros(N, np.int32)
        clab = np.zeros(N, np.int32)
        dc = np.zeros(d.size, np.int32) if want_dc_index else None
        cap = C.c_int(N)
        self._ck(self._lib.perch_depth2cloud(self._h, _ptr(d), _ptr(planes[0]), _ptr(planes[1]), _ptr(planes[2]), *args,
                                             C.byref(cap), _ptr(soa), _ptr(aos), _ptr(col), _ptr(dc), _ptr(pmap), _ptr(clab)))
        return {
            "result_cloud_eigen": aos, "result_cloud": soa, "result_cloud_color": col, "dc_index": dc,
            "rendered_cloud_point_num": N, "cloud_pose_map": pmap,
            "result_observed_cloud_label": clab if lab is not None else None,
        }

    # -- the hot path ------------------------------------------------------------------------------------------------
    def render_cuda_multi_unified(self, stage: str, poses, pose_model_map, pose_segmentation_label=None,
                                  pose_observed_points_total=None, *, stride=1, cost_type=0,
                                  calculate_observed_cost=True, sensor_resolution=0.01, color_distance_threshold=20.0,
                                  occlusion_threshold=1.0, single_result_image=0, do_icp=False,
                                  cloud_capacity: Optional[int] = None, want_dc_index=False, icp_max_iterations=0,
                                  icp_k_correspondences=0, icp_rotation_epsilon=-1.0, icp_transformation_epsilon=-1.0):
        """cuda_renderer::render_cuda_multi_unified (renderer.h:221-268) against the resident models / scene.

        stage: "RENDER" | "CLOUD" | "COST" | "DEBUG".  poses: (P,4,4) float32 mat4x4 (x100 scaled).  Returns a dict
        keyed like the reference's output arguments (result_depth, result_color, result_cloud, result_cloud_color,
        result_cloud_point_num, result_cloud_pose_map, rendered_cost, observed_cost, points_diff_cost) plus `counts`
        (3,P) int32 = rendered points, unexplained rendered points, explained observed points.  With do_icp the poses
        are first refined by batched GICP (renderer.cu:1688-1817); clouds and costs then refer to `adjusted_poses`
        (P,4,4), which is returned too.  icp_* <= 0 / < 0 select the reference's hard-coded values."""
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32).reshape(-1)
        assert len(pmm) == P, "pose_model_map length != number of poses"
        plab = None
        if pose_segmentation_label is not None and len(pose_segmentation_label) > 0:
            plab = _c(pose_segmentation_label, np.int32).reshape(-1)
            assert len(plab) == P
        ptot = None if pose_observed_points_total is None else _c(pose_observed_points_total, np.float32).reshape(-1)
        prm = perch_params(stride, cost_type, int(bool(calculate_observed_cost)), sensor_resolution, color_distance_threshold,
                           occlusion_threshold, int(single_result_image), int(bool(do_icp)), int(icp_max_iterations),
                           int(icp_k_correspondences), float(icp_rotation_epsilon), float(icp_transformation_epsilon))

        def call(out):
            return self._lib.perch_render_cost(self._h, stage.encode(), _ptr(P4), P, _ptr(pmm), _ptr(plab), _ptr(ptot),
                                               C.byref(prm), C.byref(out))

        return self._run_stage(stage, P, prm, call, cloud_capacity, want_dc_index)

    def render_cost_unfused(self, poses, pose_model_map, pose_segmentation_label=None, pose_observed_points_total=None, *,
                            stride=1, cost_type=0, calculate_observed_cost=True, sensor_resolution=0.01,
                            color_distance_threshold=20.0, occlusion_threshold=1.0):
        """perch_render_cost_unfused: the COST outputs through the unfused kernels (full raster -> cloud -> brute-force
        nearest neighbour -> counters); a device-side cross-check of the fused kernel at any batch size."""
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32).reshape(-1)
        plab = None
        if pose_segmentation_label is not None and len(pose_segmentation_label) > 0:
            plab = _c(pose_segmentation_label, np.int32).reshape(-1)
        ptot = None if pose_observed_points_total is None else _c(pose_observed_points_total, np.float32).reshape(-1)
        prm = perch_params(stride, cost_type, int(bool(calculate_observed_cost)), sensor_resolution, color_distance_threshold,
                           occlusion_threshold, 0, 0, 0, 0, -1.0, -1.0)

        def call(out):
            return self._lib.perch_render_cost_unfused(self._h, _ptr(P4), P, _ptr(pmm), _ptr(plab), _ptr(ptot), C.byref(prm),
                                                       C.byref(out))

        return self._run_stage("COST", P, prm, call)

    def _run_stage(self, stage: str, P: int, prm, call, cloud_capacity=None, want_dc_index=False):
        """Allocate the host outputs `stage` produces for P poses, run `call(perch_outputs)`, return them as a dict."""
        out = perch_outputs()
        res = {}
        if prm.do_icp:
            res["adjusted_poses"] = np.zeros((P, 4, 4), np.float32)
            out.adjusted_poses = res["adjusted_poses"].ctypes.data
        W, H = self.width, self.height
        stride, single_result_image = prm.stride, prm.single_result_image
        st = stage
        images = st in ("RENDER", "DEBUG")
        clouds = st == "DEBUG" or "CLOUD" in st
        costs = st == "DEBUG" or "COST" in st
        n_img = 1 if single_result_image else P
        if images:
            res["result_depth"] = np.zeros((n_img, H, W), np.int32)
            res["result_color"] = np.zeros((3, n_img, H, W), np.uint8)
            out.depth = res["result_depth"].ctypes.data
            out.red, out.green, out.blue = (res["result_color"][k].ctypes.data for k in range(3))
        if clouds:
            if cloud_capacity is None:
                cloud_capacity = P * (W // stride) * ((H + stride - 1) // stride)
            cap = int(cloud_capacity)
            cloud = np.zeros((3, cap), np.float32)
            ccol = np.zeros((3, cap), np.uint8)
            cmap = np.zeros(cap, np.int32)
            out.cloud, out.cloud_color, out.cloud_pose_map = cloud.ctypes.data, ccol.ctypes.data, cmap.ctypes.data
            out.cloud_capacity = cap
            if want_dc_index:
                res["result_dc_index"] = np.zeros(P * W * H, np.int32)
                out.dc_index = res["result_dc_index"].ctypes.data
        if costs:
            res["rendered_cost"] = np.zeros(P, np.float32)
            res["observed_cost"] = np.zeros(P, np.float32)
            res["points_diff_cost"] = np.zeros(P, np.float32)
            res["counts"] = np.zeros((3, P), np.int32)
            out.rendered_cost = res["rendered_cost"].ctypes.data
            out.observed_cost = res["observed_cost"].ctypes.data
            out.points_diff_cost = res["points_diff_cost"].ctypes.data
            out.counts = res["counts"].ctypes.data
        self._ck(call(out))
        if clouds:
            n = min(out.cloud_point_num, cap)
            res["result_cloud_point_num"] = out.cloud_point_num
            res["result_cloud"] = cloud[:, :n].copy()
            res["result_cloud_color"] = ccol[:, :n].copy()
            res["result_cloud_pose_map"] = cmap[:n].copy()
        return res

    # -- observed-scene preparation -------------------------------------------------------------------
    def set_input(self, depth, color=None, predicted_mask=None, *, median_blur_ksize=0, depth_factor=100.0, stride=1,
                  bounds=None, camera_transform=None, source_division_factor=0.0) -> int:
        """The GPU branch of EnvObjectRecognition::SetInput (search_env.cpp:5880-6069): depth is the decoded image,
        uint8 (3-DoF; median-blurred with `median_blur_ksize`) or uint16 (6-DoF); color = 3 planes as
        input_color_image_vec; bounds = {x_max,x_min,y_max,y_min,z_max,z_min} with camera_transform (4x4, camera optical
        frame -> world).  Leaves the source depth (depth / source_division_factor) and the observed cloud resident;
        returns observed_point_num."""
        from ._capi import perch_input
        d = np.ascontiguousarray(depth)
        assert d.dtype in (np.uint8, np.uint16), "depth must be uint8 or uint16"
        keep = [d.reshape(-1)]
        inp = perch_input()
        inp.depth, inp.depth_bits = keep[0].ctypes.data, 8 * d.dtype.itemsize
        if color is not None:
            for name, plane in zip(("red", "green", "blue"), color):
                keep.append(_c(plane, np.uint8).reshape(-1))
                setattr(inp, name, keep[-1].ctypes.data)
        if predicted_mask is not None:
            keep.append(_c(predicted_mask, np.uint8).reshape(-1))
            inp.predicted_mask = keep[-1].ctypes.data
        inp.median_blur_ksize, inp.depth_factor, inp.stride = int(median_blur_ksize), float(depth_factor), int(stride)
        if bounds is not None:
            keep.append(_c(bounds, np.float64).reshape(6))
            inp.bounds6 = keep[-1].ctypes.data
            keep.append(np.ascontiguousarray(np.asarray(camera_transform, np.float32).reshape(4, 4).T).reshape(16))
            inp.camera_transform16 = keep[-1].ctypes.data
        inp.source_division_factor = float(source_division_factor)
        n = C.c_int(0)
        self._ck(self._lib.perch_set_input(self._h, C.byref(inp), C.byref(n)))
        return n.value

    def get_observed(self):
        """The observed cloud of the last set_input, in depth2cloud_global's order: xyz (N,3), rgb (3,N), label (N,)."""
        n = C.c_int(0)
        self._ck(self._lib.perch_get_observed(self._h, None, None, None, 0, C.byref(n)))
        xyz, rgb, lab = np.zeros((n.value, 3), np.float32), np.zeros((3, n.value), np.uint8), np.zeros(n.value, np.int32)
        self._ck(self._lib.perch_get_observed(self._h, _ptr(xyz), _ptr(rgb), _ptr(lab), n.value, C.byref(n)))
        return xyz, rgb, lab

    def get_source_depth(self) -> np.ndarray:
        out = np.zeros((self.height, self.width), np.int32)
        self._ck(self._lib.perch_get_source_depth(self._h, _ptr(out)))
        return out

    def get_source_label(self):
        """The label image the occlusion rule currently tests against, or None."""
        out = np.zeros((self.height, self.width), np.uint8)
        has = C.c_int(0)
        self._ck(self._lib.perch_get_source_label(self._h, _ptr(out), C.byref(has)))
        return out if has.value else None

    # -- composition mode: placed objects rendered into the source image on the device -------------
    def compose_source(self, poses=None, pose_model_map=None, pose_segmentation_label=None, mode="merge") -> None:
        """Render the already-placed objects into the source depth the occlusion rule tests against.  mode: "replace"
        (source = placed objects only; the reference's tree flow, search_env.cpp:1718-1781, 2054-2460), "merge" (nearer
        of the observed image and the placed objects) or "none" (back to the observed image)."""
        m = {"none": _capi.PERCH_COMPOSE_NONE, "replace": _capi.PERCH_COMPOSE_REPLACE, "merge": _capi.PERCH_COMPOSE_MERGE}[mode]
        n = 0 if poses is None else len(poses)
        p = None if n == 0 else _c(poses, np.float32).reshape(n, 16)
        pm = None if n == 0 else _c(pose_model_map, np.int32)
        lab = None if (n == 0 or pose_segmentation_label is None) else _c(pose_segmentation_label, np.int32)
        self._ck(self._lib.perch_compose_source(self._h, _ptr(p), n, _ptr(pm), _ptr(lab), m))

    def tree_flags(self, poses, pose_model_map, pose_segmentation_label=None, occlusion_threshold=1.0) -> dict:
        """pose_occluded / pose_occluded_other / clutter_cost of the reference's USE_TREE / USE_CLUTTER branches
        (image_renderer.cuh:187-206, 467-482), per covered pixel on its minimum depth, + the covered-pixel count."""
        n = len(poses)
        p = _c(poses, np.float32).reshape(n, 16)
        pm = _c(pose_model_map, np.int32)
        lab = None if pose_segmentation_label is None else _c(pose_segmentation_label, np.int32)
        occ, other = np.zeros(n, np.int32), np.zeros(n, np.int32)
        clutter, total = np.zeros(n, np.float32), np.zeros(n, np.float32)
        self._ck(self._lib.perch_tree_flags(self._h, _ptr(p), n, _ptr(pm), _ptr(lab), float(occlusion_threshold), _ptr(occ),
                                            _ptr(other), _ptr(clutter), _ptr(total)))
        return {"pose_occluded": occ, "pose_occluded_other": other, "clutter_cost": clutter, "total_points": total}

    # -- caller glue: ObjectState lists (perception_b200/states.py has the host-side types) --------
    def set_env(self, env, metas) -> None:
        """EnvParams / PerchParams fields + per-model ObjectModel metadata (states.Env, list of states.ModelMeta)."""
        from . import states as S
        e = S.env_struct(env)
        m, fp = S.meta_array(metas)
        self._n_env_models = len(metas)
        self._ck(self._lib.perch_set_env(self._h, C.byref(e), C.cast(m, C.c_void_p), len(metas), _ptr(fp), len(fp)))

    def set_search_cloud(self, xyz, ranges=None, constraint_xyz=None) -> None:
        """The clouds IsValidPose / the cylinder count search (search_env.cpp:5624-5662); ranges: (R,2) [begin,end)."""
        pts = _c(xyz, np.float32).reshape(-1, 3)
        rg = None if ranges is None else _c(ranges, np.int32).reshape(-1, 2)
        cc = None if constraint_xyz is None else _c(constraint_xyz, np.float32).reshape(-1, 3)
        self._ck(self._lib.perch_set_search_cloud(self._h, _ptr(pts), len(pts), _ptr(rg), 0 if rg is None else len(rg),
                                                  _ptr(cc), 0 if cc is None else len(cc)))

    def assemble_states(self, states):
        """The per-pose loop of GetStateImagesUnifiedGPU (search_env.cpp:1555-1622): returns poses (P,4,4) float32,
        pose_model_map, pose_segmentation_label, pose_observed_points_total."""
        from . import states as S
        st = np.ascontiguousarray(states, dtype=S.STATE_DTYPE)
        P = len(st)
        poses = np.zeros((P, 4, 4), np.float32)
        pmm = np.zeros(P, np.int32)
        lab = np.zeros(P, np.int32)
        tot = np.zeros(P, np.float32)
        self._ck(self._lib.perch_assemble_states(self._h, _ptr(st), P, _ptr(poses), _ptr(pmm), _ptr(lab), _ptr(tot)))
        return poses, pmm, lab, tot

    def score_states(self, stage: str, states=None, *, stride=1, cost_type=0, calculate_observed_cost=True,
                     sensor_resolution=0.01, color_distance_threshold=20.0, occlusion_threshold=1.0, do_icp=False,
                     cloud_capacity=None, icp_max_iterations=0, icp_k_correspondences=0, icp_rotation_epsilon=-1.0,
                     icp_transformation_epsilon=-1.0):
        """GetStateImagesUnifiedGPU (search_env.cpp:1505-1717) in one call: assemble the states on the device and run
        `stage` on them.  states=None scores the list the last generate_states left on the device."""
        from . import states as S
        if states is None:
            st, P = None, self._n_generated
        else:
            st = np.ascontiguousarray(states, dtype=S.STATE_DTYPE)
            P = len(st)
        prm = perch_params(stride, cost_type, int(bool(calculate_observed_cost)), sensor_resolution, color_distance_threshold,
                           occlusion_threshold, 0, int(bool(do_icp)), int(icp_max_iterations), int(icp_k_correspondences),
                           float(icp_rotation_epsilon), float(icp_transformation_epsilon))

        def call(out):
            return self._lib.perch_score_states(self._h, stage.encode(), _ptr(st), P, C.byref(prm), C.byref(out))

        return self._run_stage(stage, P, prm, call, cloud_capacity)

    def validate_states(self, states, placed=None, after_refinement=False) -> np.ndarray:
        """EnvObjectRecognition::IsValidPose (search_env.cpp:309-528) per state; returns a bool array."""
        from . import states as S
        st = np.ascontiguousarray(states, dtype=S.STATE_DTYPE)
        pl = None if placed is None or len(placed) == 0 else np.ascontiguousarray(placed, dtype=S.STATE_DTYPE)
        valid = np.zeros(len(st), np.uint8)
        self._ck(self._lib.perch_validate_states(self._h, _ptr(st), len(st), _ptr(pl), 0 if pl is None else len(pl),
                                                 int(bool(after_refinement)), _ptr(valid)))
        return valid.astype(bool)

    def generate_states(self, placed=None, fetch=True):
        """The 3-DoF branch of GenerateSuccessorStates (search_env.cpp:7268-7472).  Returns the successor states
        (states.STATE_DTYPE array; None with fetch=False -- the list stays on the device for score_states(None))."""
        from . import states as S
        pl = None if placed is None or len(placed) == 0 else np.ascontiguousarray(placed, dtype=S.STATE_DTYPE)
        n = C.c_int(0)
        self._ck(self._lib.perch_generate_states(self._h, _ptr(pl), 0 if pl is None else len(pl), None, 0, C.byref(n)))
        self._n_generated = n.value
        if not fetch:
            return None
        out = np.zeros(n.value, S.STATE_DTYPE)
        if n.value:
            # the list is deterministic: a second pass with a buffer returns the same states
            self._ck(self._lib.perch_generate_states(self._h, _ptr(pl), 0 if pl is None else len(pl), _ptr(out), n.value,
                                                     C.byref(n)))
        return out

    @property
    def n_generated(self) -> int:
        return self._n_generated

    def gicp(self, poses, pose_model_map, pose_segmentation_label=None, *, stride=1, sensor_resolution=0.01,
             occlusion_threshold=1.0, icp_max_iterations=0, icp_k_correspondences=0, icp_rotation_epsilon=-1.0,
             icp_transformation_epsilon=-1.0):
        """Batched GICP refinement alone (perch_gicp): returns adjusted_poses (P,4,4; x100-scaled), transforms
        (P,12: R row-major then t, metres), iterations (P,), error (P,)."""
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32).reshape(-1)
        plab = None
        if pose_segmentation_label is not None and len(pose_segmentation_label) > 0:
            plab = _c(pose_segmentation_label, np.int32).reshape(-1)
        prm = perch_params(stride, 0, 0, sensor_resolution, 0.0, occlusion_threshold, 0, 1, int(icp_max_iterations),
                           int(icp_k_correspondences), float(icp_rotation_epsilon), float(icp_transformation_epsilon))
        adj = np.zeros((P, 4, 4), np.float32)
        tf = np.zeros((P, 12), np.float32)
        it = np.zeros(P, np.int32)
        err = np.zeros(P, np.float32)
        self._ck(self._lib.perch_gicp(self._h, _ptr(P4), P, _ptr(pmm), _ptr(plab), C.byref(prm), _ptr(adj), _ptr(tf),
                                      _ptr(it), _ptr(err)))
        return {"adjusted_poses": adj, "transforms": tf, "iterations": it, "error": err}

    def icp_set_scene(self, scene_depth, width, height, K, max_dist_diff=0.1, want_buffers=False):
        """Scene_projective::init_Scene_projective_cuda (depth_scene.h:19-25): depth in cm (int32), K 3x3."""
        d = _c(scene_depth, np.int32).reshape(-1)
        assert d.size == width * height
        K9 = _c(K, np.float32).reshape(9)
        pcd = np.zeros((height * width, 3), np.float32) if want_buffers else None
        nrm = np.zeros((height * width, 3), np.float32) if want_buffers else None
        self._ck(self._lib.perch_icp_set_scene(self._h, _ptr(d), width, height, _ptr(K9), max_dist_diff, _ptr(pcd), _ptr(nrm)))
        return (pcd, nrm) if want_buffers else None

    def icp_point2plane(self, clouds, relative_fitness=1e-5, relative_rmse=1e-5, max_iteration=30, nn_max_dist=None):
        """cuda_icp::ICP_Point2Plane_cuda<Scene_projective> (icp.h:92-99) for a list of (n_i,3) clouds, or, with
        nn_max_dist (reference: 0.01 m), ICP_Point2Plane_cuda<Scene_nn>: nearest-scene-point association.  Returns
        transformation (B,4,4), fitness (B,), inlier_rmse (B,), iterations (B,), and the transformed clouds."""
        clouds = [np.ascontiguousarray(c, np.float32).reshape(-1, 3) for c in clouds]
        off = np.zeros(len(clouds) + 1, np.int32)
        off[1:] = np.cumsum([len(c) for c in clouds])
        pts = np.ascontiguousarray(np.concatenate(clouds) if clouds else np.zeros((0, 3), np.float32), dtype=np.float32)
        B = len(clouds)
        T = np.zeros((B, 4, 4), np.float32)
        fit = np.zeros(B, np.float32); rmse = np.zeros(B, np.float32); it = np.zeros(B, np.int32)
        out = np.zeros_like(pts)
        if nn_max_dist is None:
            self._ck(self._lib.perch_icp_point2plane(self._h, _ptr(pts), _ptr(off), B, relative_fitness, relative_rmse,
                                                     max_iteration, _ptr(T), _ptr(fit), _ptr(rmse), _ptr(it), _ptr(out)))
        else:
            self._ck(self._lib.perch_icp_point2plane_nn(self._h, _ptr(pts), _ptr(off), B, float(nn_max_dist), relative_fitness,
                                                        relative_rmse, max_iteration, _ptr(T), _ptr(fit), _ptr(rmse), _ptr(it),
                                                        _ptr(out)))
        return {"transformation": T, "fitness": fit, "inlier_rmse": rmse, "iterations": it,
                "clouds": [out[off[i]:off[i + 1]] for i in range(B)]}

    # -- introspection ---------------------------------------------------------------------------------------------
    def last_timing(self):
        ms, n = C.c_float(0), C.c_int(0)
        self._ck(self._lib.perch_last_timing(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def stats(self):
        s = _capi.perch_gpu_stats()
        self._ck(self._lib.perch_get_stats(self._h, C.byref(s)))
        return {"icp_runtime": s.icp_runtime, "peak_memory_usage": s.peak_memory_usage}

    # -- device-resident scoring (bench / pipelining) ------------------------------------------------------------------
    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._ck(self._lib.perch_device_alloc(self._h, nbytes, C.byref(p)))
        return p.value

    def device_free(self, ptr: int) -> None:
        self._ck(self._lib.perch_device_free(self._h, C.c_void_p(ptr)))

    def h2d(self, dptr: int, arr: np.ndarray) -> None:
        a = np.ascontiguousarray(arr)
        self._ck(self._lib.perch_memcpy_h2d(self._h, C.c_void_p(dptr), _ptr(a), a.nbytes))

    def d2h(self, arr: np.ndarray, dptr: int) -> None:
        assert arr.flags["C_CONTIGUOUS"]
        self._ck(self._lib.perch_memcpy_d2h(self._h, _ptr(arr), C.c_void_p(dptr), arr.nbytes))

    def render_cost_device(self, d_poses: int, n_poses: int, d_pose_model_map: int, d_pose_label: Optional[int],
                           d_pose_obs_total: Optional[int], d_costs: int, d_counts: Optional[int], *, stride, cost_type,
                           calculate_observed_cost=True, sensor_resolution=0.01, color_distance_threshold=20.0,
                           occlusion_threshold=1.0) -> None:
        """Enqueue the fused render->cost kernel on device-resident inputs; returns without synchronising."""
        prm = perch_params(stride, cost_type, int(bool(calculate_observed_cost)), sensor_resolution, color_distance_threshold,
                           occlusion_threshold, 0, 0, 0, 0, -1.0, -1.0)
        self._ck(self._lib.perch_render_cost_device(self._h, C.c_void_p(d_poses), n_poses, C.c_void_p(d_pose_model_map),
                                                    C.c_void_p(d_pose_label) if d_pose_label else None,
                                                    C.c_void_p(d_pose_obs_total) if d_pose_obs_total else None,
                                                    C.byref(prm), C.c_void_p(d_costs),
                                                    C.c_void_p(d_counts) if d_counts else None))

    # -- a scene in one call: costs + per-model top-k reduced on the device, gathered across ranks ---------------------------
    def comm_init(self, n_ranks: int, rank: int, comm_id: bytes) -> None:
        """Attach this context to a communicator of `n_ranks` (one process per GPU): `comm_id` = comm_unique_id() of rank
        0, distributed by the launcher's own channel.  score_scene then all-gathers the per-rank top-k over NCCL."""
        assert len(comm_id) == _capi.PERCH_COMM_ID_BYTES
        buf = (C.c_char * len(comm_id)).from_buffer_copy(comm_id)
        self._ck(self._lib.perch_comm_init_rank(self._h, n_ranks, rank, C.cast(buf, C.c_void_p)))

    def comm_destroy(self) -> None:
        self._ck(self._lib.perch_comm_destroy(self._h))

    def score_scene(self, poses, pose_model_map, pose_segmentation_label=None, pose_observed_points_total=None, *,
                    n_models: int, k: int = 1, pose_id_offset: int = 0, out: Optional[dict] = None, want_costs=True,
                    stride=1, cost_type=0, calculate_observed_cost=True, sensor_resolution=0.01,
                    color_distance_threshold=20.0, occlusion_threshold=1.0):
        """perch_score_scene: this rank's pose block -> fused kernel -> device top-k -> (NCCL all-gather) -> merged
        per-model best-k of the whole scene.  Returns (hits (n_models, k) structured array, costs dict or None).
        `out` = preallocated (e.g. pinned) arrays rendered_cost / observed_cost / points_diff_cost (P,) and counts (3,P)."""
        return _score_scene(self._lib, self._lib.perch_score_scene, self._h, poses, pose_model_map, pose_segmentation_label,
                            pose_observed_points_total, n_models, k, pose_id_offset, out, want_costs,
                            perch_params(stride, cost_type, int(bool(calculate_observed_cost)), sensor_resolution,
                                         color_distance_threshold, occlusion_threshold, 0, 0, 0, 0, -1.0, -1.0))

    def synchronize(self) -> None:
        self._ck(self._lib.perch_synchronize(self._h))

    def l2_flush(self) -> None:
        self._ck(self._lib.perch_l2_flush(self._h))

    def timer_start(self) -> None:
        self._ck(self._lib.perch_timer_start(self._h))

    def timer_stop(self) -> None:
        self._ck(self._lib.perch_timer_stop(self._h))

    def timer_elapsed(self) -> float:
        ms = C.c_float(0)
        self._ck(self._lib.perch_timer_elapsed(self._h, C.byref(ms)))
        return ms.value

    def knn(self, ref, query, k: int):
        """k nearest reference points of every query point with the kNN-CUDA contract (knncuda.h:33-100).
        ref: (dim, ref_nb), query: (dim, query_nb), one row per dimension.  Returns (knn_dist, knn_index), both
        (k, query_nb): Euclidean distances ascending, 0-based indices, ties by ascending index."""
        r = np.ascontiguousarray(ref, dtype=np.float32)
        q = np.ascontiguousarray(query, dtype=np.float32)
        if r.ndim != 2 or q.ndim != 2 or r.shape[0] != q.shape[0]:
            raise ValueError("ref and query must be (dim, n) arrays of the same dim")
        dist = np.zeros((k, q.shape[1]), np.float32)
        index = np.zeros((k, q.shape[1]), np.int32)
        self._ck(self._lib.perch_knn(self._h, _ptr(r), r.shape[1], _ptr(q), q.shape[1], r.shape[0], k, _ptr(dist), _ptr(index)))
        return dist, index


HIT_DTYPE = np.dtype([("cost", "<i4"), ("pose_id", "<i4"), ("rendered_cost", "<f4"), ("observed_cost", "<f4")])


def comm_unique_id() -> bytes:
    """perch_comm_unique_id: the bytes rank 0 hands to every rank's Renderer.comm_init."""
    lib = _capi.load_library()
    buf = C.create_string_buffer(_capi.PERCH_COMM_ID_BYTES)
    _capi.check(lib, lib.perch_comm_unique_id(C.cast(buf, C.c_void_p), _capi.PERCH_COMM_ID_BYTES))
    return buf.raw


def _score_scene(lib, fn, handle, poses, pmm, plab, ptot, n_models, k, pose_id_offset, out, want_costs, prm):
    P4 = _c(poses, np.float32).reshape(-1, 16)
    P = len(P4)
    pmm = _c(pmm, np.int32).reshape(-1)
    assert len(pmm) == P, "pose_model_map length != number of poses"
    plab = None if plab is None or len(plab) == 0 else _c(plab, np.int32).reshape(-1)
    ptot = None if ptot is None else _c(ptot, np.float32).reshape(-1)
    res = None
    po = perch_outputs()
    if want_costs:
        res = out if out is not None else {"rendered_cost": np.zeros(P, np.float32), "observed_cost": np.zeros(P, np.float32),
                                           "points_diff_cost": np.zeros(P, np.float32), "counts": np.zeros((3, P), np.int32)}
        assert res["counts"].shape == (3, P) and res["rendered_cost"].shape == (P,)
        po.rendered_cost = res["rendered_cost"].ctypes.data
        po.observed_cost = res["observed_cost"].ctypes.data
        po.points_diff_cost = res["points_diff_cost"].ctypes.data
        po.counts = res["counts"].ctypes.data
    hits = np.zeros((n_models, k), dtype=HIT_DTYPE)
    args = [handle, _ptr(P4), P, _ptr(pmm), _ptr(plab), _ptr(ptot), C.byref(prm)]
    if pose_id_offset is not None:
        args.append(pose_id_offset)
    _capi.check(lib, fn(*args, n_models, k, C.byref(po) if want_costs else None, _ptr(hits)))
    return hits, res


class MultiRenderer:
    """perch_multi: one process, several GPUs -- a context, a host thread and a stream per device, the pose list cut
    into contiguous blocks, per-device top-k all-gathered over NCCL."""

    def __init__(self, devices: Sequence[int]):
        self._lib = _capi.load_library()
        dev = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        _capi.check(self._lib, self._lib.perch_create_multi(_ptr(dev), len(dev), C.byref(h)))
        self._h = h
        self.n_devices = len(dev)

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.perch_destroy_multi(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, code: int) -> None:
        _capi.check(self._lib, code)

    def set_camera(self, width, height, proj_mat, cx, cy, fx, fy, depth_factor=100.0) -> None:
        p = _c(proj_mat, np.float32).reshape(16)
        self._ck(self._lib.perch_multi_set_camera(self._h, width, height, _ptr(p), cx, cy, fx, fy, depth_factor))

    def set_models(self, tris, tris_model_count) -> None:
        t = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        self._ck(self._lib.perch_multi_set_models(self._h, _ptr(t), len(t), _ptr(cnt), len(cnt)))

    def set_scene(self, source_depth, source_color=None, source_mask_label=None) -> None:
        d = _c(source_depth, np.int32).reshape(-1)
        planes = [None] * 3 if source_color is None else [_c(source_color[k], np.uint8).reshape(-1) for k in range(3)]
        lab = None if source_mask_label is None or len(source_mask_label) == 0 else _c(source_mask_label, np.uint8).reshape(-1)
        self._ck(self._lib.perch_multi_set_scene(self._h, _ptr(d), _ptr(planes[0]), _ptr(planes[1]), _ptr(planes[2]), _ptr(lab)))

    def set_observed(self, observed_xyz, observed_color=None, observed_label=None) -> None:
        xyz = _c(observed_xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        col = None if observed_color is None else _c(observed_color, np.uint8).reshape(3, n)
        lab = None if observed_label is None else _c(observed_label, np.int32).reshape(n)
        self._ck(self._lib.perch_multi_set_observed(self._h, _ptr(xyz), _ptr(col), _ptr(lab), n))

    def score_scene(self, poses, pose_model_map, pose_segmentation_label=None, pose_observed_points_total=None, *,
                    n_models: int, k: int = 1, out: Optional[dict] = None, want_costs=True, stride=1, cost_type=0,
                    calculate_observed_cost=True, sensor_resolution=0.01, color_distance_threshold=20.0,
                    occlusion_threshold=1.0):
        """perch_multi_score on the WHOLE pose list (sharded inside the library).  Returns (hits, costs) like
        Renderer.score_scene; pose ids are indices into the list."""
        return _score_scene(self._lib, self._lib.perch_multi_score, self._h, poses, pose_model_map, pose_segmentation_label,
                            pose_observed_points_total, n_models, k, None, out, want_costs,
                            perch_params(stride, cost_type, int(bool(calculate_observed_cost)), sensor_resolution,
                                         color_distance_threshold, occlusion_threshold, 0, 0, 0, 0, -1.0, -1.0))

    def last_kernel_ms(self) -> float:
        ms = C.c_float(0)
        self._ck(self._lib.perch_multi_last_timing(self._h, C.byref(ms)))
        return ms.value


def topk(rendered_cost, observed_cost, pose_model_map, n_models: int, k: int = 1, pose_id_offset: int = 0) -> np.ndarray:
    """Per-model best-k candidates by the caller's rule (search_env.cpp:2016-2036, 2542-2583).  Returns a
    structured array (n_models, k) with fields cost, pose_id, rendered_cost, observed_cost (cost INT32_MAX = empty)."""
    lib = _capi.load_library()
    r = _c(rendered_cost, np.float32)
    o = _c(observed_cost, np.float32)
    m = _c(pose_model_map, np.int32)
    out = np.zeros((n_models, k), dtype=HIT_DTYPE)
    _capi.check(lib, lib.perch_topk(_ptr(r), _ptr(o), _ptr(m), len(r), pose_id_offset, n_models, k, _ptr(out)))
    return out


def _knn_entry(ref, ref_nb, query, query_nb, dim, k, device=0):
    r = np.ascontiguousarray(ref, dtype=np.float32).reshape(dim, ref_nb)
    q = np.ascontiguousarray(query, dtype=np.float32).reshape(dim, query_nb)
    with Renderer(device) as rr:
        return rr.knn(r, q, k)


def knn_cuda_global(ref, ref_nb, query, query_nb, dim, k, device=0):
    """cuda_renderer::knn_cuda_global (knncuda.h:33-52) minus the output pointers: returns (knn_dist, knn_index)."""
    return _knn_entry(ref, ref_nb, query, query_nb, dim, k, device)


def knn_cuda_texture(ref, ref_nb, query, query_nb, dim, k, device=0):
    """cuda_renderer::knn_cuda_texture (knncuda.h:55-74): same contract, same kernel here."""
    return _knn_entry(ref, ref_nb, query, query_nb, dim, k, device)


def knn_cublas(ref, ref_nb, query, query_nb, dim, k, device=0):
    """cuda_renderer::knn_cublas (knncuda.h:77-100): same contract, same kernel here (no GEMM: K = dim = 3 is no
    contraction worth a tensor core, and the ||r||^2+||q||^2-2rq form loses precision)."""
    return _knn_entry(ref, ref_nb, query, query_nb, dim, k, device)


# ---- stateless wrappers with the reference's full argument lists -------------------------------------------------------
def depth2cloud_global(depth_data, color_data, width, height, num_poses, pose_occluded, kCameraCX, kCameraCY, kCameraFX,
                       kCameraFY, depth_factor, stride, point_dim=3, label_mask_data=None, observed_cloud_bounds=None,
                       camera_transform=None, device=0):
    """Argument order of cuda_renderer::depth2cloud_global (renderer.h:131-153) minus the output references, which are
    returned as a dict.  `pose_occluded` and `point_dim` are accepted and ignored like in the reference."""
    with Renderer(device) as r:
        return r.depth2cloud_global(depth_data, color_data, width, height, num_poses, kCameraCX, kCameraCY, kCameraFX,
                                    kCameraFY, depth_factor, stride, label_mask_data, observed_cloud_bounds, camera_transform)


def render_cuda_multi_unified(stage, tris, poses, pose_model_map, tris_model_count, width, height, proj_mat, source_depth,
                              source_color, single_result_image, source_mask_label, pose_segmentation_label, stride,
                              point_dim, depth_factor, kCameraCX, kCameraCY, kCameraFX, kCameraFY, observed_depth_eigen,
                              observed_color, observed_point_num, pose_observed_points_total, result_observed_cloud_label,
                              cost_type, calculate_observed_cost, sensor_resolution, color_distance_threshold,
                              occlusion_threshold, do_icp=False, device=0):
    """Argument order of cuda_renderer::render_cuda_multi_unified (renderer.h:221-268) minus outputs (returned as a dict)
    and minus `observed_depth` (the SoA duplicate of observed_depth_eigen) / `clutter_cost` (USE_CLUTTER is compiled out,
    model.h:18).  Uploads everything on every call, like the reference."""
    with Renderer(device) as r:
        r.set_camera(width, height, proj_mat, kCameraCX, kCameraCY, kCameraFX, kCameraFY, depth_factor)
        r.set_models(tris, tris_model_count)
        r.set_scene(source_depth, source_color, source_mask_label)
        has_label = pose_segmentation_label is not None and len(pose_segmentation_label) > 0
        xyz = np.asarray(observed_depth_eigen, np.float32).reshape(-1, 3)[:observed_point_num]
        col = None if observed_color is None else np.asarray(observed_color, np.uint8).reshape(3, -1)[:, :observed_point_num]
        r.set_observed(xyz, col, result_observed_cloud_label if has_label else None)
        return r.render_cuda_multi_unified(
            stage, poses, pose_model_map, pose_segmentation_label, pose_observed_points_total, stride=stride,
            cost_type=cost_type, calculate_observed_cost=calculate_observed_cost, sensor_resolution=sensor_resolution,
            color_distance_threshold=color_distance_threshold, occlusion_threshold=occlusion_threshold,
            single_result_image=single_result_image, do_icp=do_icp)
