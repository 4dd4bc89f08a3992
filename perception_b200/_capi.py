"""ctypes binding of libperch_b200.so -- one Python function per C-ABI entry point in include/perch_b200.h."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libperch_b200.so"

PERCH_OK, PERCH_ERR_INVALID, PERCH_ERR_CUDA, PERCH_ERR_NO_DEVICE, PERCH_ERR_CAPACITY = range(5)


class PerchError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"perch_b200 error {code}: {msg}")
        self.code = code


class perch_params(C.Structure):
    _fields_ = [
        ("stride", C.c_int),
        ("cost_type", C.c_int),
        ("calculate_observed_cost", C.c_int),
        ("sensor_resolution", C.c_float),
        ("color_distance_threshold", C.c_float),
        ("occlusion_threshold", C.c_float),
        ("single_result_image", C.c_int),
        ("do_icp", C.c_int),
        ("icp_max_iterations", C.c_int),
        ("icp_k_correspondences", C.c_int),
        ("icp_rotation_epsilon", C.c_float),
        ("icp_transformation_epsilon", C.c_float),
    ]


class perch_outputs(C.Structure):
    _fields_ = [
        ("depth", C.c_void_p),
        ("red", C.c_void_p),
        ("green", C.c_void_p),
        ("blue", C.c_void_p),
        ("cloud", C.c_void_p),
        ("cloud_color", C.c_void_p),
        ("cloud_pose_map", C.c_void_p),
        ("dc_index", C.c_void_p),
        ("cloud_capacity", C.c_int),
        ("cloud_point_num", C.c_int),
        ("rendered_cost", C.c_void_p),
        ("observed_cost", C.c_void_p),
        ("points_diff_cost", C.c_void_p),
        ("counts", C.c_void_p),
        ("adjusted_poses", C.c_void_p),
    ]


class perch_gpu_stats(C.Structure):
    _fields_ = [("icp_runtime", C.c_float), ("peak_memory_usage", C.c_double)]


class perch_hit(C.Structure):
    _fields_ = [("cost", C.c_int32), ("pose_id", C.c_int32), ("rendered_cost", C.c_float), ("observed_cost", C.c_float)]


class perch_state(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("x", "y", "z", "roll", "pitch", "yaw", "qx", "qy", "qz", "qw")] + [
        ("model_id", C.c_int32),
        ("segmentation_label_id", C.c_int32),
    ]


class perch_model_meta(C.Structure):
    _fields_ = [
        ("preprocessing_transform", C.c_float * 16),
        ("circumscribed_radius", C.c_double),
        ("circumscribed_radius_3d", C.c_double),
        ("inscribed_radius", C.c_double),
        ("inflation_factor", C.c_double),
        ("search_resolution", C.c_double),
        ("symmetric", C.c_int32),
        ("symmetry_mode", C.c_int32),
        ("footprint_offset", C.c_int32),
        ("footprint_count", C.c_int32),
    ]


class perch_env(C.Structure):
    _fields_ = [
        ("cam_to_world", C.c_double * 16),
        ("x_min", C.c_double),
        ("x_max", C.c_double),
        ("y_min", C.c_double),
        ("y_max", C.c_double),
        ("table_height", C.c_double),
        ("res", C.c_double),
        ("theta_res", C.c_double),
        ("footprint_tolerance", C.c_double),
        ("use_external_pose_list", C.c_int32),
        ("use_cylinder_observed", C.c_int32),
        ("min_neighbor_points_for_valid_pose", C.c_int32),
        ("use_adaptive_resolution", C.c_int32),
        ("use_model_specific_search_resolution", C.c_int32),
        ("min_points_for_constraint_cloud", C.c_int32),
    ]


class perch_input(C.Structure):
    _fields_ = [
        ("depth", C.c_void_p),
        ("depth_bits", C.c_int),
        ("red", C.c_void_p),
        ("green", C.c_void_p),
        ("blue", C.c_void_p),
        ("predicted_mask", C.c_void_p),
        ("median_blur_ksize", C.c_int),
        ("depth_factor", C.c_float),
        ("stride", C.c_int),
        ("bounds6", C.c_void_p),
        ("camera_transform16", C.c_void_p),
        ("source_division_factor", C.c_float),
    ]


# name -> (restype, argtypes); the list doubles as the "every symbol of include/perch_b200.h" check in tests
_P = C.c_void_p
_V = C.c_void_p
SIGNATURES = {
    "perch_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "perch_destroy": (None, [_P]),
    "perch_last_error": (C.c_char_p, []),
    "perch_abi_version": (C.c_int, []),
    "perch_get_stats": (C.c_int, [_P, C.POINTER(perch_gpu_stats)]),
    "perch_last_timing": (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "perch_gicp": (C.c_int, [_P, _V, C.c_int, _V, _V, C.POINTER(perch_params), _V, _V, _V, _V]),
    "perch_icp_set_scene": (C.c_int, [_P, _V, C.c_int, C.c_int, _V, C.c_float, _V, _V]),
    "perch_icp_point2plane": (C.c_int, [_P, _V, _V, C.c_int, C.c_float, C.c_float, C.c_int, _V, _V, _V, _V, _V]),
    "perch_icp_point2plane_nn": (C.c_int, [_P, _V, _V, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, _V, _V, _V, _V, _V]),
    "perch_icp_get_scene_device": (C.c_int, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "perch_icp_point2plane_device": (C.c_int, [_P, _V, C.c_int, _V, _V, C.c_int, C.c_int, _V, C.c_float, C.c_int, C.c_float,
                                               C.c_float, C.c_int, _V, _V, _V, _V]),
    "perch_icp_depth2cloud_device": (C.c_int, [_P, _V, C.c_int, C.c_int, C.c_int, _V, C.c_int, C.c_int, C.c_int, _V, C.c_int,
                                               C.POINTER(C.c_int)]),
    "perch_knn": (C.c_int, [_P, _V, C.c_int, _V, C.c_int, C.c_int, C.c_int, _V, _V]),
    "perch_compute_proj": (C.c_int, [_P, C.c_int, C.c_int, C.c_float, C.c_float, _P]),
    "perch_set_camera": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]),
    "perch_set_models": (C.c_int, [_P, _P, C.c_int64, _P, C.c_int]),
    "perch_set_scene": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "perch_set_observed": (C.c_int, [_P, _P, _P, _P, C.c_int]),
    "perch_depth2cloud": (
        C.c_int,
        [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int,
         _P, _P, _P, C.POINTER(C.c_int), _P, _P, _P, _P, _P, _P],
    ),
    "perch_depth2cloud_static": (
        C.c_int,
        [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
         _P, _P, _P, _P, C.POINTER(C.c_int)],
    ),
    "perch_render_cost": (C.c_int, [_P, C.c_char_p, _P, C.c_int, _P, _P, _P, C.POINTER(perch_params), C.POINTER(perch_outputs)]),
    "perch_render_cost_unfused": (C.c_int, [_P, _P, C.c_int, _P, _P, _P, C.POINTER(perch_params), C.POINTER(perch_outputs)]),
    "perch_score_scene": (C.c_int, [_P, _P, C.c_int, _P, _P, _P, C.POINTER(perch_params), C.c_int, C.c_int, C.c_int,
                                    C.POINTER(perch_outputs), _P]),
    "perch_comm_unique_id": (C.c_int, [_P, C.c_size_t]),
    "perch_comm_init_rank": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "perch_comm_destroy": (C.c_int, [_P]),
    "perch_create_multi": (C.c_int, [_P, C.c_int, C.POINTER(_P)]),
    "perch_destroy_multi": (None, [_P]),
    "perch_multi_size": (C.c_int, [_P]),
    "perch_multi_context": (_P, [_P, C.c_int]),
    "perch_multi_set_camera": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]),
    "perch_multi_set_models": (C.c_int, [_P, _P, C.c_int64, _P, C.c_int]),
    "perch_multi_set_scene": (C.c_int, [_P, _P, _P, _P, _P, _P]),
    "perch_multi_set_observed": (C.c_int, [_P, _P, _P, _P, C.c_int]),
    "perch_multi_score": (C.c_int, [_P, _P, C.c_int, _P, _P, _P, C.POINTER(perch_params), C.c_int, C.c_int,
                                    C.POINTER(perch_outputs), _P]),
    "perch_multi_last_timing": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "perch_render_cost_device": (C.c_int, [_P, _P, C.c_int, _P, _P, _P, C.POINTER(perch_params), _P, _P]),
    "perch_device_alloc": (C.c_int, [_P, C.c_size_t, C.POINTER(_P)]),
    "perch_device_free": (C.c_int, [_P, _P]),
    "perch_memcpy_h2d": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "perch_memcpy_d2h": (C.c_int, [_P, _P, _P, C.c_size_t]),
    "perch_synchronize": (C.c_int, [_P]),
    "perch_l2_flush": (C.c_int, [_P]),
    "perch_timer_start": (C.c_int, [_P]),
    "perch_timer_stop": (C.c_int, [_P]),
    "perch_timer_elapsed": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "perch_topk": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _P]),
    "perch_set_env": (C.c_int, [_P, C.POINTER(perch_env), _V, C.c_int, _V, C.c_int]),
    "perch_set_search_cloud": (C.c_int, [_P, _V, C.c_int, _V, C.c_int, _V, C.c_int]),
    "perch_assemble_states": (C.c_int, [_P, _V, C.c_int, _V, _V, _V, _V]),
    "perch_score_states": (C.c_int, [_P, C.c_char_p, _V, C.c_int, C.POINTER(perch_params), C.POINTER(perch_outputs)]),
    "perch_validate_states": (C.c_int, [_P, _V, C.c_int, _V, C.c_int, C.c_int, _V]),
    "perch_generate_states": (C.c_int, [_P, _V, C.c_int, _V, C.c_int, C.POINTER(C.c_int)]),
    "perch_parse_pose_list": (C.c_int, [C.c_char_p, C.c_size_t, C.c_int, C.c_int, _V, C.c_int, C.POINTER(C.c_int)]),
    "perch_set_input": (C.c_int, [_P, C.POINTER(perch_input), C.POINTER(C.c_int)]),
    "perch_get_observed": (C.c_int, [_P, _V, _V, _V, C.c_int, C.POINTER(C.c_int)]),
    "perch_get_source_depth": (C.c_int, [_P, _V]),
    "perch_cluster_stats": (C.c_int, [_V, C.c_int64, _V, C.c_int, _V]),
    "perch_get_source_label": (C.c_int, [_P, _V, C.POINTER(C.c_int)]),
    "perch_compose_source": (C.c_int, [_P, _V, C.c_int, _V, _V, C.c_int]),
    "perch_tree_flags": (C.c_int, [_P, _V, C.c_int, _V, _V, C.c_float, _V, _V, _V, _V]),
}

PERCH_COMPOSE_NONE, PERCH_COMPOSE_REPLACE, PERCH_COMPOSE_MERGE = range(3)
PERCH_COMM_ID_BYTES = 128

_lib = None


def _preload_nccl() -> None:
    """libperch_b200.so links libnccl.so.2 (the top-k all-gather of a sharded pose list).  In a process that also imports
    PyTorch, the NCCL that PyTorch bundles must be the one behind that soname -- libtorch_cuda.so needs symbols the
    system copy may lack -- whichever of the two libraries is loaded first; so load PyTorch's copy first when there is one.
    Without PyTorch the system libnccl.so.2 is found by the dynamic linker as usual."""
    import sys
    for base in sys.path:
        cand = Path(base) / "nvidia" / "nccl" / "lib" / "libnccl.so.2"
        if cand.exists():
            try:
                C.CDLL(str(cand), mode=C.RTLD_GLOBAL)
            except OSError:
                pass
            return


def load_library(path: os.PathLike | None = None) -> C.CDLL:
    """Load libperch_b200.so (built in-tree by ``make`` / ``__graft_entry__.build()``).  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    # PERCH_B200_LIB: another build of the same library (tools/: A/B runs of kernel variants, the counters build)
    p = Path(path) if path else Path(os.environ.get("PERCH_B200_LIB", LIB_PATH))
    if not p.exists():
        raise ImportError(
            f"{p} not found: build it with `make` (nvcc, sm_100a). perception_b200 has no CPU fallback."
        )
    _preload_nccl()
    lib = C.CDLL(str(p))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here == the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def check(lib: C.CDLL, code: int) -> None:
    if code != PERCH_OK:
        raise PerchError(code, lib.perch_last_error().decode("utf-8", "replace"))
