"""The C++ drop-in backend (perception_b200/csrc/compat/cuda_renderer_compat.cpp, compiled against the reference's OWN
cuda_renderer/renderer.h) called the way sbpl_perception::EnvObjectRecognition calls it (std::vector / Eigen / cv::Mat
arguments, callee-malloc'd outputs), through the extern "C" trampolines of oracle/compat_driver.cpp.  Results must equal
the oracle's: this is the drop-in boundary exercised end to end."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import Oracle
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt

pytestmark = pytest.mark.gpu
SO = Path(__file__).resolve().parents[1] / "oracle" / "_ref" / "libcuda_renderer_compat.so"

if has_gpu():
    from perception_b200 import synth
    from perception_b200.renderer import TRIANGLE_DTYPE


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def lib():
    if not SO.exists():
        pytest.skip("oracle/_ref/libcuda_renderer_compat.so not built (needs /root/reference at build time)")
    l = C.CDLL(str(SO))
    l.compat_last_error.restype = C.c_char_p
    return l


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def test_compute_proj(lib, orc):
    K = synth.camera_K().astype(np.float32)
    out = np.zeros(16, np.float32)
    assert lib.compat_compute_proj(_p(np.ascontiguousarray(K)), 640, 480, _p(out)) == 0, lib.compat_last_error()
    assert np.array_equal(out.reshape(4, 4), np.asarray(orc.compute_proj(K, 640, 480)).reshape(4, 4))


@pytest.mark.parametrize("mode", ["label", "bounds"])
def test_depth2cloud_global(lib, orc, mode):
    case = make_case(600, 3, 42)
    s = case.scene
    stride = 4
    kw = {"label_mask": s.source_label} if mode == "label" else {
        "bounds": synth.table_bounds(s), "cam_tf": np.linalg.inv(s.T_cam_table)}
    o = orc.depth2cloud(s.source_depth, s.source_color, case.width, case.height, 1, case.cx, case.cy, case.fx, case.fy,
                        100.0, stride, **kw)
    cap = int(o["n"]) + 16
    soa = np.zeros((3, cap), np.float32); aos = np.zeros((cap, 3), np.float32)
    rgb = np.zeros((3, cap), np.uint8); lab = np.zeros(cap, np.int32); n = C.c_int(0)
    rgb_in = np.ascontiguousarray(np.stack([s.source_color[k] for k in range(3)]).astype(np.uint8))
    mask = np.ascontiguousarray(s.source_label.astype(np.uint8)) if mode == "label" else None
    bounds = np.ascontiguousarray(np.asarray(kw["bounds"], np.float64)) if mode == "bounds" else None
    tf = np.ascontiguousarray(np.asarray(kw["cam_tf"], np.float32).T) if mode == "bounds" else None  # column-major
    rc = lib.compat_depth2cloud(_p(np.ascontiguousarray(s.source_depth, np.int32)), _p(rgb_in), case.width, case.height,
                                C.c_float(case.cx), C.c_float(case.cy), C.c_float(case.fx), C.c_float(case.fy),
                                C.c_float(100.0), stride, _p(mask), _p(bounds), _p(tf), cap, _p(soa), _p(aos), _p(rgb),
                                _p(lab), C.byref(n))
    assert rc == 0, lib.compat_last_error()
    assert n.value == o["n"] > 0
    assert np.array_equal(soa[:, :n.value], o["soa"][:, :n.value])
    assert np.array_equal(aos[:n.value], o["aos"][:n.value])
    assert np.array_equal(rgb[:, :n.value], o["rgb"][:, :n.value])
    if mode == "label":
        assert np.array_equal(lab[:n.value], o["label"][:n.value])


def _render(lib, stage, case, poses, pmm, lab, obs, tot, stride, cost_type, res, dof6, want_images, cloud_cap=0):
    s = case.scene
    P = len(poses)
    tris = np.ascontiguousarray(s.tris, dtype=TRIANGLE_DTYPE)
    cnt = np.ascontiguousarray(s.tris_model_count, np.int32)
    rgb_in = np.ascontiguousarray(np.stack([s.source_color[k] for k in range(3)]).astype(np.uint8))
    n_px = case.width * case.height
    depth = np.zeros(P * n_px, np.int32) if want_images else None
    rgb = np.zeros(3 * P * n_px, np.uint8) if want_images else None
    cloud = np.zeros((3, cloud_cap), np.float32) if cloud_cap else None
    cloud_n = C.c_int(0)
    costs = np.full(3 * P, np.nan, np.float32)
    rc = lib.compat_render(
        stage.encode(), _p(tris), len(tris), _p(np.ascontiguousarray(poses, np.float32)), P,
        _p(np.ascontiguousarray(pmm, np.int32)), _p(cnt), len(cnt), case.width, case.height,
        _p(np.ascontiguousarray(case.proj, np.float32)), _p(np.ascontiguousarray(s.source_depth, np.int32)), _p(rgb_in),
        _p(np.ascontiguousarray(s.source_label.astype(np.uint8))) if dof6 else None,
        _p(np.ascontiguousarray(lab, np.int32)) if dof6 else None, stride, 100,
        C.c_float(case.cx), C.c_float(case.cy), C.c_float(case.fx), C.c_float(case.fy),
        _p(np.ascontiguousarray(obs["aos"], np.float32)) if obs else None,
        _p(np.ascontiguousarray(obs["rgb"], np.uint8)) if obs else None, len(obs["aos"]) if obs else 0,
        _p(np.ascontiguousarray(tot, np.float32)) if tot is not None else None,
        _p(np.ascontiguousarray(obs["label"], np.int32)) if (obs and dof6) else None, cost_type, 1, C.c_float(res),
        C.c_float(20.0), C.c_float(1.0), _p(depth), _p(rgb), cloud_cap, _p(cloud), C.byref(cloud_n), _p(costs))
    assert rc == 0, lib.compat_last_error()
    return depth, rgb, cloud, cloud_n.value, costs.reshape(3, P)


@pytest.mark.parametrize("dof6", [False, True])
def test_render_and_cost_through_reference_signature(lib, orc, dof6):
    case = make_case(600, 3, 42)
    s = case.scene
    stride = 4
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    poses, pmm, lab = poses_near_gt(case, 10)
    if dof6:
        tot = np.bincount(obs["label"], minlength=len(s.gt_poses)).astype(np.float32)[lab]
    else:
        tot = synth.cylinder_counts(s, obs["aos"], poses, pmm)
    res, ct = (0.01, 2) if dof6 else (0.0075, 0)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          s.source_label if dof6 else None, lab if dof6 else None, 1.0)
    o = orc.costs(d_o, c_o, case.width, case.height, case.cx, case.cy, case.fx, case.fy, 100.0, stride,
                  lab if dof6 else None, obs["aos"], obs["rgb"], obs["label"] if dof6 else None, tot, ct, True, res, 20.0)
    P, n_px = len(poses), case.width * case.height
    # RENDER: images as the caller receives them (result_depth + three result_color planes)
    depth, rgb, _, _, _ = _render(lib, "RENDER", case, poses, pmm, lab, None, None, stride, ct, res, dof6, True)
    assert np.array_equal(depth.reshape(P, case.height, case.width), d_o)
    assert np.array_equal(rgb.reshape(3, P, case.height, case.width), c_o)
    # COST: the three callee-malloc'd arrays
    _, _, _, _, costs = _render(lib, "COST", case, poses, pmm, lab, obs, tot, stride, ct, res, dof6, False)
    assert np.array_equal(costs[0], o["rendered_cost"])
    assert np.array_equal(costs[1], o["observed_cost"], equal_nan=True)
    assert np.array_equal(costs[2], o["points_diff_cost"])
    # CLOUD: compacted SoA cloud with row stride = point count
    _, _, cloud, cloud_n, _ = _render(lib, "CLOUD", case, poses, pmm, lab, None, None, stride, ct, res, dof6, False,
                                      cloud_cap=P * 4096)
    oc = orc.depth2cloud(d_o.reshape(-1), None, case.width, case.height, P, case.cx, case.cy, case.fx, case.fy, 100.0, stride)
    assert cloud_n == oc["n"] > 0
    assert np.array_equal(cloud[:, :cloud_n], oc["soa"][:, :cloud_n])
