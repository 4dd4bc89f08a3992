"""Observed-scene preparation: perch_set_input = the GPU branch of SetInput
(search_env.cpp:5880-6069) + the source-depth rescale of ComputeGreedyRenderPoses (2487-2498) in one call, everything
resident.  OpenCV is absent: cv::medianBlur is restated (exact k x k median, replicated border) and pinned here against
scipy.ndimage; the cloud kernels are the ones already pinned against the reference's compute_point_clouds.cuh."""
import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import Oracle, oracle_divide_depth, oracle_median_blur_u8
from perception_b200 import synth
from scenes import make_case, poses_near_gt

gpu = pytest.mark.gpu


def _noisy_u8(h=97, w=123, seed=0):
    rng = np.random.default_rng(seed)
    img = (60 + 40 * np.sin(np.arange(w) / 9.0)[None] + 30 * np.cos(np.arange(h) / 7.0)[:, None]).astype(np.int32)
    img += rng.integers(-20, 20, size=img.shape)
    img[rng.uniform(size=img.shape) < 0.1] = 0          # depth holes, what the blur is there to remove
    return np.clip(img, 0, 255).astype(np.uint8)


@pytest.mark.parametrize("k", [1, 3, 5, 9, 17])
def test_oracle_median_matches_scipy(k):
    from scipy import ndimage
    img = _noisy_u8()
    ref = ndimage.median_filter(img, size=k, mode="nearest") if k > 1 else img
    assert np.array_equal(oracle_median_blur_u8(img, k), ref)


def test_oracle_divide_depth():
    d = np.array([0, 1, 99, 100, 101, 12345, 65535, 7], np.int32)
    assert np.array_equal(oracle_divide_depth(d, 100.0), (d.astype(np.float32) / np.float32(100.0)).astype(np.int32))
    assert np.array_equal(oracle_divide_depth(d, 1.0), d)


def _scene_u8(case):
    """NDDS-style 8-bit depth image of the test scene (values are the renderer's cm, which fit a byte here)."""
    d = case.scene.source_depth
    assert d.max() < 256
    return d.astype(np.uint8)


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
@pytest.mark.parametrize("k", [0, 5, 17])
def test_set_input_3dof_equals_manual_path(k):
    from perception_b200.renderer import Renderer
    case = make_case()
    s = case.scene
    orc = Oracle()
    depth8 = _scene_u8(case)
    T_tc = np.linalg.inv(s.T_cam_table)
    bounds = synth.table_bounds(s)
    stride = 4
    blurred = oracle_median_blur_u8(depth8, k) if k > 1 else depth8
    ref_cloud = orc.depth2cloud(blurred.astype(np.int32), s.source_color, case.width, case.height, 1, case.cx, case.cy,
                                case.fx, case.fy, 100.0, stride, bounds=bounds, cam_tf=T_tc)
    poses, pmm, _ = poses_near_gt(case, 24)
    tot = np.full(len(poses), float(ref_cloud["n"]), np.float32)
    kw = dict(stride=stride, cost_type=0, sensor_resolution=0.0075)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, 100.0)
        r.set_models(s.tris, s.tris_model_count)
        n = r.set_input(depth8, s.source_color, median_blur_ksize=k, depth_factor=100.0, stride=stride, bounds=bounds,
                        camera_transform=T_tc)
        xyz, rgb, lab = r.get_observed()
        assert n == ref_cloud["n"] == len(xyz) > 100
        assert np.array_equal(xyz, ref_cloud["aos"]) and np.array_equal(rgb, ref_cloud["rgb"]) and np.all(lab == -1)
        assert np.array_equal(r.get_source_depth(), blurred.astype(np.int32))
        a = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, **kw)
        # the same scene set up by hand through the three separate entry points
        r.set_scene(blurred.astype(np.int32), s.source_color)
        r.set_observed(ref_cloud["aos"], ref_cloud["rgb"])
        b = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, **kw)
        for key in ("rendered_cost", "observed_cost", "points_diff_cost", "counts"):
            assert np.array_equal(a[key], b[key], equal_nan=True), key


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_set_input_6dof_equals_manual_path():
    from perception_b200.renderer import Renderer
    case = make_case()
    s = case.scene
    orc = Oracle()
    # YCB-style 16-bit depth: 10000 units per metre, the renderer works in cm => division factor 100
    rng = np.random.default_rng(3)
    depth16 = (s.source_depth * 100 + rng.integers(0, 100, size=s.source_depth.shape) * (s.source_depth > 0)).astype(np.uint16)
    stride = 4
    ref_cloud = orc.depth2cloud(depth16.astype(np.int32), s.source_color, case.width, case.height, 1, case.cx, case.cy,
                                case.fx, case.fy, 10000.0, stride, label_mask=s.source_label)
    src = oracle_divide_depth(depth16.astype(np.int32), 100.0)
    assert np.array_equal(src, s.source_depth)
    poses, pmm, plab = poses_near_gt(case, 24)
    seg = np.bincount(ref_cloud["label"], minlength=len(s.gt_poses)).astype(np.float32)
    tot = seg[plab]
    kw = dict(stride=stride, cost_type=2, sensor_resolution=0.01)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, 100.0)
        r.set_models(s.tris, s.tris_model_count)
        n = r.set_input(depth16, s.source_color, s.source_label, depth_factor=10000.0, stride=stride,
                        source_division_factor=100.0)
        xyz, rgb, lab = r.get_observed()
        assert n == ref_cloud["n"] > 100
        assert np.array_equal(xyz, ref_cloud["aos"]) and np.array_equal(rgb, ref_cloud["rgb"])
        assert np.array_equal(lab, ref_cloud["label"])
        assert np.array_equal(r.get_source_depth(), src)
        a = r.render_cuda_multi_unified("COST", poses, pmm, plab, tot, **kw)
        r.set_scene(src, s.source_color, s.source_label)
        r.set_observed(ref_cloud["aos"], ref_cloud["rgb"], ref_cloud["label"])
        b = r.render_cuda_multi_unified("COST", poses, pmm, plab, tot, **kw)
        for key in ("rendered_cost", "observed_cost", "points_diff_cost", "counts"):
            assert np.array_equal(a[key], b[key], equal_nan=True), key
        assert a["counts"][0].max() > 0


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_median_blur_kernel_sizes_and_errors():
    from perception_b200.renderer import Renderer, compute_proj
    img = _noisy_u8(48, 64, seed=5)
    K = np.array([[60, 0, 32], [0, 60, 24], [0, 0, 1]], np.float32)
    with Renderer(0) as r:
        r.set_camera(64, 48, compute_proj(K, 64, 48), 32, 24, 60, 60, 100.0)
        for k in (3, 7, 31):
            r.set_input(img, median_blur_ksize=k, stride=2)
            assert np.array_equal(r.get_source_depth(), oracle_median_blur_u8(img, k).astype(np.int32)), k
        with pytest.raises(Exception, match="odd"):
            r.set_input(img, median_blur_ksize=4, stride=2)
        with pytest.raises(Exception, match="stride"):
            r.set_input(img, stride=5)
        with pytest.raises(Exception, match="8-bit"):
            r.set_input(img.astype(np.uint16), median_blur_ksize=3, stride=2)


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_scene_buffers_follow_the_camera_size():
    """A context that held a SMALL scene and is then given a larger camera must not reuse the small scene buffers:
    perch_set_input used to copy the new image into whatever allocation was there (device memory overrun).  The large
    scene set after the small one must score exactly like the same scene on a fresh context, and going back to the small
    size must work too."""
    from perception_b200.renderer import Renderer
    small, big = make_case(300, 3, 7, 160, 120), make_case()
    sb = big.scene
    depth8 = _scene_u8(big)
    T_tc = np.linalg.inv(sb.T_cam_table)
    bounds = synth.table_bounds(sb)
    poses, pmm, _ = poses_near_gt(big, 16)
    kw = dict(stride=4, cost_type=0, sensor_resolution=0.0075)

    def big_scene(r):
        r.set_camera(big.width, big.height, big.proj, big.cx, big.cy, big.fx, big.fy, 100.0)
        r.set_models(sb.tris, sb.tris_model_count)
        n = r.set_input(depth8, sb.source_color, depth_factor=100.0, stride=4, bounds=bounds, camera_transform=T_tc)
        return r.render_cuda_multi_unified("COST", poses, pmm, None, np.full(len(poses), float(n), np.float32), **kw)

    with Renderer(0) as r:
        ss = small.scene
        r.set_camera(small.width, small.height, small.proj, small.cx, small.cy, small.fx, small.fy, 100.0)
        r.set_models(ss.tris, ss.tris_model_count)
        r.set_scene(ss.source_depth, ss.source_color, ss.source_label)   # 160 x 120 buffers, labelled
        after_small = big_scene(r)                                       # 640 x 480, unlabelled, through perch_set_input
        r.set_camera(small.width, small.height, small.proj, small.cx, small.cy, small.fx, small.fy, 100.0)
        r.set_scene(ss.source_depth, ss.source_color, None)
        assert r.get_source_depth().shape == (small.height, small.width)
    with Renderer(0) as r:
        fresh = big_scene(r)
    for key in ("rendered_cost", "observed_cost", "points_diff_cost", "counts"):
        assert np.array_equal(after_small[key], fresh[key], equal_nan=True), key
    assert (fresh["counts"][0] > 0).any()
