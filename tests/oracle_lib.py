"""ctypes access to the TEST oracles: oracle/_build/libperch_oracle.so (CPU restatement) and
oracle/_ref/libperch_ref.so (the reference's own CUDA kernels, GPU box only).  Test infrastructure only."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from perception_b200.renderer import TRIANGLE_DTYPE

ROOT = Path(__file__).resolve().parent.parent
ORACLE_SO = ROOT / "oracle" / "_build" / "libperch_oracle.so"
REF_SO = ROOT / "oracle" / "_ref" / "libperch_ref.so"
REF_TREE_SO = ROOT / "oracle" / "_ref" / "libperch_ref_tree.so"

_vp = C.c_void_p


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(str(ORACLE_SO))
        self.lib.oracle_depth2cloud.restype = C.c_int
        self.lib.oracle_color_distance_rgb.restype = C.c_double
        self.lib.oracle_color_distance_rgb.argtypes = [C.c_uint8] * 6

    def compute_proj(self, K, w, h, near=10.0, far=10000.0):
        out = np.zeros(16, np.float32)
        Kc = _c(K, np.float32).reshape(9)
        self.lib.oracle_compute_proj(_p(Kc), C.c_int(w), C.c_int(h), C.c_float(near), C.c_float(far), _p(out))
        return out.reshape(4, 4)

    def render(self, tris, tris_model_count, poses, pose_model_map, width, height, proj, source_depth=None,
               source_label=None, pose_label=None, occlusion_threshold=1.0, single_result_image=0):
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32)
        proj = _c(proj, np.float32).reshape(16)
        sd = np.zeros(width * height, np.int32) if source_depth is None else _c(source_depth, np.int32).reshape(-1)
        sl = None if source_label is None else _c(source_label, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        n_img = 1 if single_result_image else P
        depth = np.zeros((n_img, height, width), np.int32)
        rgb = np.zeros((3, n_img, height, width), np.uint8)
        self.lib.oracle_render(_p(tris), C.c_int(len(tris)), _p(P4), C.c_int(P), _p(pmm), _p(cnt), C.c_int(len(cnt)),
                               C.c_int(width), C.c_int(height), _p(proj), _p(sd), _p(sl), _p(pl),
                               C.c_float(occlusion_threshold), C.c_int(single_result_image), _p(depth), _p(rgb[0]),
                               _p(rgb[1]), _p(rgb[2]))
        return depth, rgb

    def render_tree(self, tris, tris_model_count, poses, pose_model_map, width, height, proj, source_depth=None,
                    source_label=None, pose_label=None, occlusion_threshold=1.0):
        """oracle_render + the USE_TREE / USE_CLUTTER vectors in both readings (serial triangle order, canonical)."""
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32)
        proj = _c(proj, np.float32).reshape(16)
        sd = np.zeros(width * height, np.int32) if source_depth is None else _c(source_depth, np.int32).reshape(-1)
        sl = None if source_label is None else _c(source_label, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        depth = np.zeros((P, height, width), np.int32)
        rgb = np.zeros((3, P, height, width), np.uint8)
        out = {}
        for kind in ("serial", "canon"):
            out[kind] = {"pose_occluded": np.zeros(P, np.int32), "pose_occluded_other": np.zeros(P, np.int32),
                         "clutter_points": np.zeros(P, np.float32), "total_points": np.zeros(P, np.float32)}
        vecs = [_p(out[k][n]) for k in ("serial", "canon")
                for n in ("pose_occluded", "pose_occluded_other", "clutter_points", "total_points")]
        self.lib.oracle_render_tree(_p(tris), C.c_int(len(tris)), _p(P4), C.c_int(P), _p(pmm), _p(cnt), C.c_int(len(cnt)),
                                    C.c_int(width), C.c_int(height), _p(proj), _p(sd), _p(sl), _p(pl),
                                    C.c_float(occlusion_threshold), _p(depth), _p(rgb[0]), _p(rgb[1]), _p(rgb[2]), *vecs)
        for k in out:
            with np.errstate(invalid="ignore", divide="ignore"):
                # thrust::divides, then multiplies by 100 (image_renderer.cuh:467-482), in float32
                out[k]["clutter_cost"] = (out[k]["clutter_points"] / out[k]["total_points"]) * np.float32(100)
        out["depth"] = depth
        return out

    def depth2cloud(self, depth, color, width, height, num_images, cx, cy, fx, fy, depth_factor, stride, label_mask=None,
                    pose_label=None, bounds=None, cam_tf=None, want_dc_index=False):
        d = _c(depth, np.int32).reshape(-1)
        planes = [None] * 3 if color is None else [_c(color[k], np.uint8).reshape(-1) for k in range(3)]
        lm = None if label_mask is None else _c(label_mask, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        bd = None if bounds is None else _c(bounds, np.float64)
        tf = None if cam_tf is None else np.ascontiguousarray(np.asarray(cam_tf, np.float32).T).reshape(16)
        args = (_p(d), _p(planes[0]), _p(planes[1]), _p(planes[2]), C.c_int(width), C.c_int(height), C.c_int(num_images),
                C.c_float(cx), C.c_float(cy), C.c_float(fx), C.c_float(fy), C.c_float(depth_factor), C.c_int(stride),
                _p(lm), _p(pl), _p(bd), _p(tf))
        n = self.lib.oracle_depth2cloud(*args, None, None, None, None, None, None, C.c_int(0))
        soa = np.zeros((3, n), np.float32)
        aos = np.zeros((n, 3), np.float32)
        rgb = np.zeros((3, n), np.uint8)
        pmap = np.zeros(n, np.int32)
        lab = np.zeros(n, np.int32)
        dc = np.zeros(d.size, np.int32) if want_dc_index else None
        self.lib.oracle_depth2cloud(*args, _p(soa), _p(aos), _p(rgb), _p(pmap), _p(lab), _p(dc), C.c_int(n))
        return {"n": n, "soa": soa, "aos": aos, "rgb": rgb, "pose_map": pmap, "label": lab, "dc_index": dc}

    def knn_k1(self, src, tgt, src_label=None, label_begin=None, label_end=None):
        s = _c(src, np.float32).reshape(-1, 3)
        t = _c(tgt, np.float32).reshape(-1, 3)
        sl = None if src_label is None else _c(src_label, np.int32)
        lb = None if label_begin is None else _c(label_begin, np.int32)
        le = None if label_end is None else _c(label_end, np.int32)
        dist = np.zeros(len(s), np.float32)
        idx = np.zeros(len(s), np.int32)
        self.lib.oracle_knn_k1(_p(s), C.c_int(len(s)), _p(t), C.c_int(len(t)), _p(sl), _p(lb), _p(le),
                               C.c_int(0 if lb is None else len(lb)), _p(dist), _p(idx))
        return dist, idx

    def costs(self, depth, color, width, height, cx, cy, fx, fy, depth_factor, stride, pose_label, obs_xyz, obs_rgb,
              obs_label, pose_obs_total, cost_type, calc_observed, sensor_resolution, color_distance_threshold,
              color_eps=1e-3):
        d = _c(depth, np.int32).reshape(-1, height, width)
        P = len(d)
        planes = [_c(color[k], np.uint8).reshape(-1) for k in range(3)]
        pl = None if pose_label is None else _c(pose_label, np.int32)
        ox = _c(obs_xyz, np.float32).reshape(-1, 3)
        n_obs = len(ox)
        orgb = np.zeros((3, n_obs), np.uint8) if obs_rgb is None else _c(obs_rgb, np.uint8).reshape(3, n_obs)
        ol = None if obs_label is None else _c(obs_label, np.int32)
        tot = None if pose_obs_total is None else _c(pose_obs_total, np.float32)
        rc = np.zeros(P, np.float32)
        oc = np.zeros(P, np.float32)
        dc = np.zeros(P, np.float32)
        counts = np.zeros((3, P), np.int32)
        near = np.zeros(P, np.int32)
        self.lib.oracle_costs(_p(d), _p(planes[0]), _p(planes[1]), _p(planes[2]), C.c_int(P), C.c_int(width), C.c_int(height),
                              C.c_float(cx), C.c_float(cy), C.c_float(fx), C.c_float(fy), C.c_float(depth_factor),
                              C.c_int(stride), _p(pl), _p(ox), _p(orgb), _p(ol), C.c_int(n_obs), _p(tot), C.c_int(cost_type),
                              C.c_int(int(calc_observed)), C.c_float(sensor_resolution), C.c_float(color_distance_threshold),
                              _p(rc), _p(oc), _p(dc), _p(counts), _p(near), C.c_double(color_eps))
        return {"rendered_cost": rc, "observed_cost": oc, "points_diff_cost": dc, "counts": counts, "near_threshold": near}

    def score_poses(self, tris, tris_model_count, poses, pose_model_map, width, height, proj, source_depth, source_label,
                    pose_label, cx, cy, fx, fy, depth_factor, stride, obs_xyz, obs_rgb, obs_label, pose_obs_total, cost_type,
                    calc_observed, sensor_resolution, color_distance_threshold, occlusion_threshold=1.0):
        """oracle_score_poses: render + cost per pose (OpenMP over poses), no images kept -- full-size parity checks."""
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32)
        proj = _c(proj, np.float32).reshape(16)
        sd = _c(source_depth, np.int32).reshape(-1)
        sl = None if source_label is None else _c(source_label, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        ox = _c(obs_xyz, np.float32).reshape(-1, 3)
        n_obs = len(ox)
        orgb = np.zeros((3, n_obs), np.uint8) if obs_rgb is None else _c(obs_rgb, np.uint8).reshape(3, n_obs)
        ol = None if obs_label is None else _c(obs_label, np.int32)
        tot = None if pose_obs_total is None else _c(pose_obs_total, np.float32)
        rc, oc, dc = np.zeros(P, np.float32), np.zeros(P, np.float32), np.zeros(P, np.float32)
        counts = np.zeros((3, P), np.int32)
        self.lib.oracle_score_poses(_p(tris), C.c_int(len(tris)), _p(P4), C.c_int(P), _p(pmm), _p(cnt), C.c_int(len(cnt)),
                                    C.c_int(width), C.c_int(height), _p(proj), _p(sd), _p(sl), _p(pl),
                                    C.c_float(occlusion_threshold), C.c_float(cx), C.c_float(cy), C.c_float(fx), C.c_float(fy),
                                    C.c_float(depth_factor), C.c_int(stride), _p(ox), _p(orgb), _p(ol), C.c_int(n_obs), _p(tot),
                                    C.c_int(cost_type), C.c_int(int(calc_observed)), C.c_float(sensor_resolution),
                                    C.c_float(color_distance_threshold), _p(rc), _p(oc), _p(dc), _p(counts))
        return {"rendered_cost": rc, "observed_cost": oc, "points_diff_cost": dc, "counts": counts}

    def color_distance_rgb(self, c1, c2):
        return self.lib.oracle_color_distance_rgb(*[int(v) for v in c1], *[int(v) for v in c2])


    def knn(self, ref, query, k, fused=True):
        """oracle_knn: kNN-CUDA contract; fused=True = device arithmetic (fma chain, root after selection),
        fused=False = knn_c's host arithmetic (root before selection)."""
        r = _c(ref, np.float32)
        q = _c(query, np.float32)
        dist = np.zeros((k, q.shape[1]), np.float32)
        idx = np.zeros((k, q.shape[1]), np.int32)
        rc = self.lib.oracle_knn(_p(r), C.c_int(r.shape[1]), _p(q), C.c_int(q.shape[1]), C.c_int(r.shape[0]), C.c_int(k),
                                 C.c_int(1 if fused else 0), _p(dist), _p(idx))
        assert rc == 0, rc
        return dist, idx


    def gicp(self, src, tgt, tgt_label=None, label=0, k=10, max_iterations=150, rotation_epsilon=2e-3,
             translation_epsilon=5e-4):
        """oracle_gicp (double precision): returns (T12 = R row-major + t, iterations, error)."""
        s = _c(src, np.float32).reshape(-1, 3)
        t = _c(tgt, np.float32).reshape(-1, 3)
        tl = None if tgt_label is None else _c(tgt_label, np.int32)
        T = np.zeros(12, np.float64)
        err = C.c_double(0)
        self.lib.oracle_gicp.restype = C.c_int
        it = self.lib.oracle_gicp(_p(s), C.c_int(len(s)), _p(t), C.c_int(len(t)), _p(tl), C.c_int(label), C.c_int(k),
                                  C.c_int(max_iterations), C.c_double(rotation_epsilon), C.c_double(translation_epsilon),
                                  _p(T), C.byref(err))
        return T, it, err.value


    def icp_scene(self, depth, W, H, K):
        d = _c(depth, np.int32).reshape(-1)
        K9 = _c(K, np.float32).reshape(9)
        pcd = np.zeros((W * H, 3), np.float32)
        nrm = np.zeros((W * H, 3), np.float32)
        self.lib.oracle_icp_scene(_p(d), C.c_int(W), C.c_int(H), _p(K9), _p(pcd), _p(nrm))
        return pcd, nrm

    def icp_point2plane(self, cloud, scene_pcd, scene_normals, W, H, K, max_dist_diff=0.1, relative_fitness=1e-5,
                        relative_rmse=1e-5, max_iteration=30, nn=False):
        """nn=False: Scene_projective association; nn=True: Scene_nn (nearest scene point within max_dist_diff)."""
        pts = np.array(cloud, np.float32, copy=True).reshape(-1, 3)
        K9 = _c(K, np.float32).reshape(9)
        T = np.zeros(16, np.float32)
        fit, rmse = C.c_float(0), C.c_float(0)
        fn = self.lib.oracle_icp_point2plane_nn if nn else self.lib.oracle_icp_point2plane
        fn.restype = C.c_int
        it = fn(_p(pts), C.c_int(len(pts)), _p(_c(scene_pcd, np.float32)),
                                             _p(_c(scene_normals, np.float32)), C.c_int(W), C.c_int(H), _p(K9),
                                             C.c_float(max_dist_diff), C.c_float(relative_fitness), C.c_float(relative_rmse),
                                             C.c_int(max_iteration), _p(T), C.byref(fit), C.byref(rmse))
        return {"transformation": T.reshape(4, 4), "fitness": fit.value, "inlier_rmse": rmse.value, "iterations": it, "cloud": pts}


ICP_REF_SO = ROOT / "oracle" / "_ref" / "libicp_ref.so"


class ReferenceIcp:
    """The reference's cuda_icp package compiled from its own sources (oracle/ref_icp_harness.cu).  Scene preparation
    is host code (runs without a GPU); the ICP loop needs one."""

    def __init__(self):
        if not ICP_REF_SO.exists():
            raise FileNotFoundError(f"{ICP_REF_SO} missing: build with `make -C oracle ref` where /root/reference exists")
        self.lib = C.CDLL(str(ICP_REF_SO))

    def scene(self, depth, W, H, K):
        d = _c(depth, np.int32).reshape(-1)
        K9 = _c(K, np.float32).reshape(9)
        pcd = np.zeros((W * H, 3), np.float32)
        nrm = np.zeros((W * H, 3), np.float32)
        assert self.lib.ref_icp_scene(_p(d), C.c_int(W), C.c_int(H), _p(K9), _p(pcd), _p(nrm)) == 0
        return pcd, nrm

    def point2plane(self, cloud, scene_pcd, scene_normals, W, H, K, max_dist_diff=0.1, relative_fitness=1e-5,
                    relative_rmse=1e-5, max_iteration=30):
        pts = np.array(cloud, np.float32, copy=True).reshape(-1, 3)
        K9 = _c(K, np.float32).reshape(9)
        T = np.zeros(16, np.float32)
        fit, rmse = C.c_float(0), C.c_float(0)
        rc = self.lib.ref_icp_point2plane(_p(pts), C.c_int(len(pts)), _p(_c(scene_pcd, np.float32)),
                                          _p(_c(scene_normals, np.float32)), C.c_int(W), C.c_int(H), _p(K9),
                                          C.c_float(max_dist_diff), C.c_float(relative_fitness), C.c_float(relative_rmse),
                                          C.c_int(max_iteration), _p(T), C.byref(fit), C.byref(rmse))
        assert rc == 0
        return {"transformation": T.reshape(4, 4), "fitness": fit.value, "inlier_rmse": rmse.value, "cloud": pts}


    def point2plane_nn(self, cloud, depth, W, H, K, relative_fitness=1e-5, relative_rmse=1e-5, max_iteration=30):
        """ICP_Point2Plane_cuda<Scene_nn>: scene cloud, normals and kd-tree built by the reference's init_Scene_nn_cuda."""
        pts = np.array(cloud, np.float32, copy=True).reshape(-1, 3)
        d = _c(depth, np.int32).reshape(-1)
        K9 = _c(K, np.float32).reshape(9)
        T = np.zeros(16, np.float32)
        fit, rmse = C.c_float(0), C.c_float(0)
        rc = self.lib.ref_icp_nn_point2plane(_p(pts), C.c_int(len(pts)), _p(d), C.c_int(W), C.c_int(H), _p(K9),
                                             C.c_float(relative_fitness), C.c_float(relative_rmse), C.c_int(max_iteration),
                                             _p(T), C.byref(fit), C.byref(rmse))
        assert rc == 0
        return {"transformation": T.reshape(4, 4), "fitness": fit.value, "inlier_rmse": rmse.value, "cloud": pts}


KNN_REF_SO = ROOT / "oracle" / "_ref" / "libknn_ref.so"


class ReferenceKnn:
    """The reference's own kNN package compiled unmodified from cuda_renderer/src/cuda/knncuda.cu (C++ mangled names).
    knn_c is host code (runs without a GPU); the three device variants need one."""
    SYMS = {"knn_c": "_ZN13cuda_renderer5knn_cEPKfiS1_iiiPfPi",
            "knn_cuda_global": "_ZN13cuda_renderer15knn_cuda_globalEPKfiS1_iiiPfPi",
            "knn_cuda_texture": "_ZN13cuda_renderer16knn_cuda_textureEPKfiS1_iiiPfPi",
            "knn_cublas": "_ZN13cuda_renderer10knn_cublasEPKfiS1_iiiPfPi"}

    def __init__(self):
        if not KNN_REF_SO.exists():
            raise FileNotFoundError(f"{KNN_REF_SO} missing: build with `make -C oracle ref` where /root/reference exists")
        self.lib = C.CDLL(str(KNN_REF_SO))

    def run(self, which, ref, query, k):
        fn = getattr(self.lib, self.SYMS[which])
        fn.restype = C.c_bool
        r = _c(ref, np.float32)
        q = _c(query, np.float32)
        dist = np.zeros((k, q.shape[1]), np.float32)
        idx = np.zeros((k, q.shape[1]), np.int32)
        # the reference printf's progress/errors on stdout: keep the test output clean
        ok = fn(_p(r), C.c_int(r.shape[1]), _p(q), C.c_int(q.shape[1]), C.c_int(r.shape[0]), C.c_int(k), _p(dist), _p(idx))
        return bool(ok), dist, idx


class Reference:
    """The reference's own CUDA kernels (oracle/ref_harness.cu).  Needs a GPU.  tree=True loads the build with the
    reference's USE_TREE / USE_CLUTTER switches on (render_tree)."""

    def __init__(self, tree=False):
        so = REF_TREE_SO if tree else REF_SO
        if not so.exists():
            raise FileNotFoundError(f"{so} missing: build with `make -C oracle ref` where /root/reference exists")
        self.lib = C.CDLL(str(so))
        if tree:
            assert self.lib.ref_tree_enabled() == 1

    def render_tree(self, tris, tris_model_count, poses, pose_model_map, width, height, proj, source_depth=None,
                    source_label=None, pose_label=None, occlusion_threshold=1.0):
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32)
        proj = _c(proj, np.float32).reshape(16)
        sd = np.zeros(width * height, np.int32) if source_depth is None else _c(source_depth, np.int32).reshape(-1)
        sl = None if source_label is None else _c(source_label, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        depth = np.zeros((P, height, width), np.int32)
        occ, other = np.zeros(P, np.int32), np.zeros(P, np.int32)
        clutter, total = np.zeros(P, np.float32), np.zeros(P, np.float32)
        rc = self.lib.ref_render_tree(_p(tris), C.c_int(len(tris)), _p(P4), C.c_int(P), _p(pmm), _p(cnt), C.c_int(len(cnt)),
                                      C.c_int(width), C.c_int(height), _p(proj), _p(sd), _p(sl), _p(pl),
                                      C.c_float(occlusion_threshold), _p(depth), _p(occ), _p(other), _p(clutter), _p(total))
        assert rc == 0, "ref_render_tree failed"
        return {"depth": depth, "pose_occluded": occ, "pose_occluded_other": other, "clutter_cost": clutter,
                "total_points": total}

    def render(self, tris, tris_model_count, poses, pose_model_map, width, height, proj, source_depth=None,
               source_color=None, source_label=None, pose_label=None, occlusion_threshold=1.0, single_result_image=0):
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32)
        proj = _c(proj, np.float32).reshape(16)
        npx = width * height
        sd = np.zeros(npx, np.int32) if source_depth is None else _c(source_depth, np.int32).reshape(-1)
        sc = [np.zeros(npx, np.uint8)] * 3 if source_color is None else [_c(source_color[k], np.uint8).reshape(-1) for k in range(3)]
        sl = None if source_label is None else _c(source_label, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        n_img = 1 if single_result_image else P
        depth = np.zeros((n_img, height, width), np.int32)
        rgb = np.zeros((3, n_img, height, width), np.uint8)
        rc = self.lib.ref_render(_p(tris), C.c_int(len(tris)), _p(P4), C.c_int(P), _p(pmm), _p(cnt), C.c_int(len(cnt)),
                                 C.c_int(width), C.c_int(height), _p(proj), _p(sd), _p(sc[0]), _p(sc[1]), _p(sc[2]), _p(sl),
                                 _p(pl), C.c_float(occlusion_threshold), C.c_int(single_result_image), _p(depth), _p(rgb[0]),
                                 _p(rgb[1]), _p(rgb[2]))
        assert rc == 0, "ref_render failed"
        return depth, rgb

    def depth2cloud(self, depth, color, width, height, num_images, cx, cy, fx, fy, depth_factor, stride, label_mask=None,
                    bounds=None, cam_tf=None, want_dc_index=False):
        d = _c(depth, np.int32).reshape(-1)
        planes = [np.zeros(d.size, np.uint8)] * 3 if color is None else [_c(color[k], np.uint8).reshape(-1) for k in range(3)]
        lm = None if label_mask is None else _c(label_mask, np.uint8).reshape(-1)
        bd = None if bounds is None else _c(bounds, np.float64)
        tf = None if cam_tf is None else np.ascontiguousarray(np.asarray(cam_tf, np.float32).T).reshape(16)
        n = C.c_int(0)
        args = (_p(d), _p(planes[0]), _p(planes[1]), _p(planes[2]), C.c_int(width), C.c_int(height), C.c_int(num_images),
                C.c_float(cx), C.c_float(cy), C.c_float(fx), C.c_float(fy), C.c_float(depth_factor), C.c_int(stride),
                _p(lm), _p(bd), _p(tf))
        rc = self.lib.ref_depth2cloud(*args, C.byref(n), None, None, None, None, None, None)
        assert rc == 0
        N = n.value
        soa = np.zeros((3, N), np.float32)
        aos = np.zeros((N, 3), np.float32)
        rgb = np.zeros((3, N), np.uint8)
        pmap = np.zeros(N, np.int32)
        lab = np.zeros(N, np.int32)
        dc = np.zeros(d.size, np.int32) if want_dc_index else None
        rc = self.lib.ref_depth2cloud(*args, C.byref(n), _p(soa), _p(aos), _p(rgb), _p(pmap), _p(lab), _p(dc))
        assert rc == 0
        return {"n": N, "soa": soa, "aos": aos, "rgb": rgb, "pose_map": pmap, "label": lab, "dc_index": dc}

    def render_cost(self, tris, tris_model_count, poses, pose_model_map, width, height, proj, source_depth, source_color,
                    source_label, pose_label, stride, depth_factor, cx, cy, fx, fy, obs_xyz, obs_rgb, obs_label,
                    pose_obs_total, cost_type, calc_observed, sensor_resolution, color_distance_threshold,
                    occlusion_threshold, want_cloud=False):
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE_DTYPE)
        cnt = _c(tris_model_count, np.int32)
        P4 = _c(poses, np.float32).reshape(-1, 16)
        P = len(P4)
        pmm = _c(pose_model_map, np.int32)
        proj = _c(proj, np.float32).reshape(16)
        npx = width * height
        sd = _c(source_depth, np.int32).reshape(-1)
        sc = [np.zeros(npx, np.uint8)] * 3 if source_color is None else [_c(source_color[k], np.uint8).reshape(-1) for k in range(3)]
        sl = None if source_label is None else _c(source_label, np.uint8).reshape(-1)
        pl = None if pose_label is None else _c(pose_label, np.int32)
        ox = _c(obs_xyz, np.float32).reshape(-1, 3)
        n_obs = len(ox)
        orgb = np.zeros((3, n_obs), np.uint8) if obs_rgb is None else _c(obs_rgb, np.uint8).reshape(3, n_obs)
        ol = None if obs_label is None else _c(obs_label, np.int32)
        tot = np.zeros(P, np.float32) if pose_obs_total is None else _c(pose_obs_total, np.float32)
        rc_ = np.zeros(P, np.float32)
        oc = np.zeros(P, np.float32)
        dc = np.zeros(P, np.float32)
        n_rend = C.c_int(0)
        cap = P * (width // stride) * ((height + stride - 1) // stride) if want_cloud else 0
        cloud = np.zeros((3, max(cap, 1)), np.float32)
        pmap = np.zeros(max(cap, 1), np.int32)
        nnd = np.zeros(max(cap, 1), np.float32)
        nni = np.zeros(max(cap, 1), np.int32)
        ms = np.zeros(4, np.float32)
        rc = self.lib.ref_render_cost(
            _p(tris), C.c_int(len(tris)), _p(P4), C.c_int(P), _p(pmm), _p(cnt), C.c_int(len(cnt)), C.c_int(width),
            C.c_int(height), _p(proj), _p(sd), _p(sc[0]), _p(sc[1]), _p(sc[2]), _p(sl), _p(pl), C.c_int(stride),
            C.c_int(int(depth_factor)), C.c_float(cx), C.c_float(cy), C.c_float(fx), C.c_float(fy), _p(ox), _p(orgb), _p(ol),
            C.c_int(n_obs), _p(tot), C.c_int(cost_type), C.c_int(int(calc_observed)), C.c_float(sensor_resolution),
            C.c_float(color_distance_threshold), C.c_float(occlusion_threshold), _p(rc_), _p(oc), _p(dc), C.byref(n_rend),
            _p(cloud) if want_cloud else None, _p(pmap) if want_cloud else None, C.c_int(cap),
            _p(nnd) if want_cloud else None, _p(nni) if want_cloud else None, _p(ms))
        assert rc == 0, "ref_render_cost failed"
        out = {"rendered_cost": rc_, "observed_cost": oc, "points_diff_cost": dc, "n_rend": n_rend.value, "ms": ms}
        if want_cloud:
            n = n_rend.value
            out.update(cloud=cloud[:, :n], pose_map=pmap[:n], nn_sqdist=nnd[:n], nn_index=nni[:n])
        return out

    def cost_from_images(self, depth, color, width, height, cx, cy, fx, fy, depth_factor, stride, pose_label, obs_xyz,
                         obs_rgb, obs_label, pose_obs_total, cost_type, calc_observed, sensor_resolution,
                         color_distance_threshold):
        """Reference compute_point_clouds + (restated k=1 search) + compute_costs on given rendered images."""
        d = _c(depth, np.int32).reshape(-1, height, width)
        P = len(d)
        planes = [_c(color[k], np.uint8).reshape(-1) for k in range(3)]
        pl = None if pose_label is None else _c(pose_label, np.int32)
        ox = _c(obs_xyz, np.float32).reshape(-1, 3)
        n_obs = len(ox)
        orgb = np.zeros((3, n_obs), np.uint8) if obs_rgb is None else _c(obs_rgb, np.uint8).reshape(3, n_obs)
        ol = None if obs_label is None else _c(obs_label, np.int32)
        tot = np.zeros(P, np.float32) if pose_obs_total is None else _c(pose_obs_total, np.float32)
        rc_ = np.zeros(P, np.float32)
        oc = np.zeros(P, np.float32)
        dc = np.zeros(P, np.float32)
        n_rend = C.c_int(0)
        rc = self.lib.ref_cost_from_images(
            _p(d), _p(planes[0]), _p(planes[1]), _p(planes[2]), C.c_int(P), C.c_int(width), C.c_int(height), _p(pl),
            C.c_int(stride), C.c_int(int(depth_factor)), C.c_float(cx), C.c_float(cy), C.c_float(fx), C.c_float(fy), _p(ox),
            _p(orgb), _p(ol), C.c_int(n_obs), _p(tot), C.c_int(cost_type), C.c_int(int(calc_observed)),
            C.c_float(sensor_resolution), C.c_float(color_distance_threshold), _p(rc_), _p(oc), _p(dc), C.byref(n_rend))
        assert rc == 0, "ref_cost_from_images failed"
        return {"rendered_cost": rc_, "observed_cost": oc, "points_diff_cost": dc, "n_rend": n_rend.value}


class StatesOracle:
    """oracle/perch_states_oracle.c: CPU restatement of the caller glue (search_env.cpp:309-528, 1505-1622, 7268-7472)."""

    def __init__(self, env, metas, search_xyz=None, ranges=None, constraint_xyz=None):
        from perception_b200 import states as S
        self.S = S
        self.lib = C.CDLL(str(ORACLE_SO))
        self.lib.oracle_generate_states.restype = C.c_longlong
        self.lib.oracle_assemble_states.restype = C.c_int
        self.lib.oracle_validate_states.restype = C.c_int
        self.env = S.env_struct(env)
        self.metas, self.fp = S.meta_array(metas)
        self.n_models = len(metas)
        self.xyz = np.zeros((0, 3), np.float32) if search_xyz is None else _c(search_xyz, np.float32).reshape(-1, 3)
        self.ranges = (np.array([[0, len(self.xyz)]], np.int32) if ranges is None else _c(ranges, np.int32).reshape(-1, 2))
        self.constraint = None if constraint_xyz is None else _c(constraint_xyz, np.float32).reshape(-1, 3)

    def camera_matrix(self):
        out = np.zeros(16, np.float64)
        self.lib.oracle_camera_matrix(C.byref(self.env, type(self.env).cam_to_world.offset), _p(out))
        return out.reshape(4, 4)

    def assemble(self, states, seg_counts=None, observed_point_num=0, num_pixels=640 * 480):
        st = np.ascontiguousarray(states, dtype=self.S.STATE_DTYPE)
        P = len(st)
        sc = np.zeros(1, np.float32) if seg_counts is None else _c(seg_counts, np.float32)
        poses = np.zeros((P, 4, 4), np.float32)
        pmm = np.zeros(P, np.int32)
        lab = np.zeros(P, np.int32)
        tot = np.zeros(P, np.float32)
        rc = self.lib.oracle_assemble_states(C.byref(self.env), self.metas, C.c_int(self.n_models), _p(st), C.c_int(P), _p(sc),
                                             C.c_int(0 if seg_counts is None else len(sc)), C.c_float(observed_point_num),
                                             C.c_int(num_pixels), _p(self.xyz), _p(self.ranges), C.c_int(len(self.ranges)),
                                             _p(poses), _p(pmm), _p(lab), _p(tot))
        assert rc == 0, rc
        return poses, pmm, lab, tot

    def _search_args(self):
        nc = 0 if self.constraint is None else len(self.constraint)
        return (_p(self.fp), _p(self.xyz), _p(self.ranges), C.c_int(len(self.ranges)), _p(self.constraint), C.c_int(nc))

    def validate(self, states, placed=None, after_refinement=False):
        st = np.ascontiguousarray(states, dtype=self.S.STATE_DTYPE)
        pl = None if placed is None or len(placed) == 0 else np.ascontiguousarray(placed, dtype=self.S.STATE_DTYPE)
        valid = np.zeros(len(st), np.uint8)
        rc = self.lib.oracle_validate_states(C.byref(self.env), self.metas, C.c_int(self.n_models), *self._search_args(),
                                             _p(st), C.c_int(len(st)), _p(pl), C.c_int(0 if pl is None else len(pl)),
                                             C.c_int(int(after_refinement)), _p(valid))
        assert rc == 0, rc
        return valid.astype(bool)

    def generate(self, placed=None):
        pl = None if placed is None or len(placed) == 0 else np.ascontiguousarray(placed, dtype=self.S.STATE_DTYPE)
        args = (C.byref(self.env), self.metas, C.c_int(self.n_models), *self._search_args(), _p(pl),
                C.c_int(0 if pl is None else len(pl)))
        n = self.lib.oracle_generate_states(*args, None, C.c_longlong(0))
        assert n >= 0
        out = np.zeros(n, self.S.STATE_DTYPE)
        self.lib.oracle_generate_states(*args, _p(out), C.c_longlong(n))
        return out


def oracle_median_blur_u8(img, k):
    """oracle_median_blur_u8 (oracle/perch_states_oracle.c): cv::medianBlur restated for CV_8UC1."""
    lib = C.CDLL(str(ORACLE_SO))
    a = np.ascontiguousarray(img, np.uint8)
    out = np.zeros_like(a)
    lib.oracle_median_blur_u8(_p(a), _p(out), C.c_int(a.shape[1]), C.c_int(a.shape[0]), C.c_int(k))
    return out


def oracle_divide_depth(depth, factor):
    lib = C.CDLL(str(ORACLE_SO))
    a = np.ascontiguousarray(depth, np.int32)
    out = np.zeros_like(a)
    lib.oracle_divide_depth(_p(a), C.c_float(factor), _p(out), C.c_int(a.size))
    return out
