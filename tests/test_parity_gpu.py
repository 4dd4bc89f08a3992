"""GPU parity tests: libperch_b200.so (through its C-ABI) vs the CPU oracle and vs the reference's own CUDA kernels
(oracle/_ref/libperch_ref.so) on identical seeded inputs.

Parity classes: bit-exact int32 depth / masks / point counts; float costs equal (they are exact
functions of the counts); colour cost flip-tolerant near the threshold.  The reference rasterizer is racy
(image_renderer.cuh:161-186), so reference-vs-canonical mismatches are counted and bounded, not hidden.
"""
import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import REF_SO, Oracle, Reference
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt

pytestmark = pytest.mark.gpu

if has_gpu():
    from perception_b200 import synth
    from perception_b200.renderer import Renderer, TRIANGLE_DTYPE, make_triangles, pose_to_mat4x4, topk


@pytest.fixture(scope="module")
def orc():
    return Oracle()


@pytest.fixture(scope="module")
def ref():
    if not REF_SO.exists():
        pytest.skip("oracle/_ref/libperch_ref.so not built (needs /root/reference at build time)")
    return Reference()


def make_renderer(case, with_label=False):
    s = case.scene
    r = Renderer(0)
    r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
    r.set_models(s.tris, s.tris_model_count)
    r.set_scene(s.source_depth, s.source_color, s.source_label if with_label else None)
    return r


# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("size,n_tris", [((640, 480), 600), ((640, 480), 5000), ((160, 120), 300)])
def test_render_3dof_bit_exact(orc, ref, size, n_tris):
    case = make_case(n_tris, 3, 42, *size)
    s = case.scene
    poses, pmm, _ = poses_near_gt(case, 24)
    with make_renderer(case) as r:
        out = r.render_cuda_multi_unified("RENDER", poses, pmm, occlusion_threshold=1.0)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          None, None, 1.0)
    assert np.array_equal(out["result_depth"], d_o), f"{np.count_nonzero(out['result_depth'] != d_o)} depth pixels differ from the oracle"
    assert np.array_equal(out["result_color"], c_o)
    assert (d_o > 0).sum() > 1000
    d_r, c_r = ref.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          s.source_color, None, None, 1.0)
    _check_reference_race_budget(d_r, d_o)


def _check_reference_race_budget(d_r, d_o, max_frac=0.15):
    """The reference rasterizer is not deterministic on sm_70+ (unfenced spin lock + racy occlusion step, see
    tests/make_golden_inputs.py): report how far it strays from the canonical result and bound it.  Deviations can only
    make a pixel emptier or farther (a lost z-test / a clobbered min), never nearer than the true minimum."""
    obj = int((d_o > 0).sum())
    bad = int(np.count_nonzero(d_r != d_o))
    cleared = int(np.count_nonzero((d_r == 0) & (d_o > 0)))
    farther = int(np.count_nonzero((d_r > d_o) & (d_o > 0)))
    nearer = int(np.count_nonzero((d_r < d_o) & (d_r > 0)))
    extra = int(np.count_nonzero((d_r > 0) & (d_o == 0)))
    print(f"reference-vs-canonical: {bad} px differ of {obj} object px (cleared {cleared}, farther {farther}, nearer {nearer}, extra {extra})")
    assert bad <= max_frac * obj + 16
    assert nearer == 0, "the reference produced a depth NEARER than the min over fragments: arithmetic mismatch"


def test_render_6dof_labels_bit_exact(orc, ref):
    case = make_case(2000, 4, 7)
    s = case.scene
    poses, pmm, lab = poses_near_gt(case, 16)
    with make_renderer(case, with_label=True) as r:
        out = r.render_cuda_multi_unified("RENDER", poses, pmm, lab)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          s.source_label, lab, 1.0)
    assert np.array_equal(out["result_depth"], d_o)
    assert np.array_equal(out["result_color"], c_o)
    d_r, _ = ref.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                        s.source_color, s.source_label, lab, 1.0)
    _check_reference_race_budget(d_r, d_o)


def test_reference_kernels_exact_on_race_free_cases(orc, ref):
    """One fragment per pixel => the reference kernel is deterministic: product == oracle == reference, bit for bit."""
    import make_golden_inputs as gi
    case, tris, counts, poses, pmm = gi.single_tri_inputs()
    with Renderer(0) as r:
        r.set_camera(gi.W, gi.H, case.proj, case.cx, case.cy, case.fx, case.fy, 100.0)
        r.set_models(tris, counts)
        r.set_scene(np.zeros(gi.W * gi.H, np.int32))
        out = r.render_cuda_multi_unified("RENDER", poses, pmm)
        d_o, c_o = orc.render(tris, counts, poses, pmm, gi.W, gi.H, case.proj)
        d_r, c_r = ref.render(tris, counts, poses, pmm, gi.W, gi.H, case.proj)
        assert (d_o > 0).sum() > 10000
        assert np.array_equal(out["result_depth"], d_o) and np.array_equal(d_r, d_o)
        assert np.array_equal(out["result_color"], c_o) and np.array_equal(c_r, c_o)
        case, tris, src, lab, pose = gi.quad_inputs()
        r.set_models(tris, [2])
        for thr in (0.0, 0.5, 1.0, 2.5):
            r.set_scene(src)
            a = r.render_cuda_multi_unified("RENDER", pose, [0], occlusion_threshold=thr)
            d_o, c_o = orc.render(tris, [2], pose, [0], gi.W, gi.H, case.proj, src, None, None, thr)
            d_r, c_r = ref.render(tris, [2], pose, [0], gi.W, gi.H, case.proj, src, None, None, None, thr)
            assert np.array_equal(a["result_depth"], d_o) and np.array_equal(d_r, d_o), thr
            assert np.array_equal(a["result_color"], c_o) and np.array_equal(c_r, c_o), thr
        r.set_scene(src, None, lab)
        for pl in (0, 1):
            a = r.render_cuda_multi_unified("RENDER", pose, [0], [pl])
            d_o, c_o = orc.render(tris, [2], pose, [0], gi.W, gi.H, case.proj, src, lab, [pl], 1.0)
            d_r, c_r = ref.render(tris, [2], pose, [0], gi.W, gi.H, case.proj, src, None, lab, [pl], 1.0)
            assert np.array_equal(a["result_depth"], d_o) and np.array_equal(d_r, d_o), pl
            assert np.array_equal(a["result_color"], c_o) and np.array_equal(c_r, c_o), pl


def test_render_edge_cases(orc):
    case = make_case(600, 3, 42)
    s = case.scene
    T = s.gt_poses[0].copy()
    poses = []
    for dx, dz in [(0.0, 0.0), (0.55, 0.0), (-0.6, 0.0), (5.0, 0.0), (0.0, -0.55), (0.0, 3.0)]:
        Tt = T.copy()
        Tt[0, 3] += dx
        Tt[2, 3] += dz
        poses.append(pose_to_mat4x4(Tt))
    # box (12 huge triangles) close to the camera, and behind it
    box = len(s.tris_model_count) - 1
    Tb = np.eye(4); Tb[:3, 3] = (0.02, 0.01, 0.25)
    poses.append(pose_to_mat4x4(Tb))
    Tb2 = np.eye(4); Tb2[:3, 3] = (0.0, 0.0, -0.5)
    poses.append(pose_to_mat4x4(Tb2))
    poses = np.stack(poses)
    pmm = np.array([s.gt_models[0]] * 6 + [box, box], np.int32)
    with make_renderer(case) as r:
        out = r.render_cuda_multi_unified("RENDER", poses, pmm)
        empty = r.render_cuda_multi_unified("RENDER", np.zeros((0, 4, 4), np.float32), np.zeros(0, np.int32))
    assert empty["result_depth"].shape[0] == 0
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          None, None, 1.0)
    for i in range(len(poses)):
        assert np.array_equal(out["result_depth"][i], d_o[i]), f"pose {i}: {np.count_nonzero(out['result_depth'][i] != d_o[i])} px differ"
    assert np.array_equal(out["result_color"], c_o)
    assert (d_o[3] > 0).sum() == 0  # far off screen


def test_single_result_image(orc, ref):
    case = make_case(600, 3, 42)
    s = case.scene
    poses = np.stack([pose_to_mat4x4(T) for T in s.gt_poses])
    pmm = np.array(s.gt_models, np.int32)
    with make_renderer(case) as r:
        out = r.render_cuda_multi_unified("RENDER", poses, pmm, single_result_image=1, occlusion_threshold=100.0)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          None, None, 100.0, single_result_image=1)
    assert out["result_depth"].shape == (1, case.height, case.width)
    assert np.array_equal(out["result_depth"], d_o)
    assert np.array_equal(out["result_color"], c_o)
    d_r, _ = ref.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                        s.source_color, None, None, 100.0, single_result_image=1)
    _check_reference_race_budget(d_r, d_o)


# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["plain", "label", "bounds"])
@pytest.mark.parametrize("stride", [1, 4, 8])
def test_depth2cloud_parity(orc, ref, mode, stride):
    case = make_case(600, 3, 42)
    s = case.scene
    kw = {}
    if mode == "label":
        kw["label_mask"] = s.source_label
    if mode == "bounds":
        kw["bounds"] = synth.table_bounds(s)
        kw["cam_tf"] = np.linalg.inv(s.T_cam_table)
    o = orc.depth2cloud(s.source_depth, s.source_color, case.width, case.height, 1, case.cx, case.cy, case.fx, case.fy,
                        100.0, stride, want_dc_index=True, **kw)
    q = ref.depth2cloud(s.source_depth, s.source_color, case.width, case.height, 1, case.cx, case.cy, case.fx, case.fy,
                        100.0, stride, want_dc_index=True, **kw)
    with Renderer(0) as r:
        p = r.depth2cloud_global(s.source_depth, s.source_color, case.width, case.height, 1, case.cx, case.cy, case.fx,
                                 case.fy, 100.0, stride, kw.get("label_mask"), kw.get("bounds"), kw.get("cam_tf"),
                                 want_dc_index=True)
    assert p["rendered_cloud_point_num"] == o["n"] == q["n"] > 0
    for name, a, b, c in [("soa", p["result_cloud"], o["soa"], q["soa"]), ("aos", p["result_cloud_eigen"], o["aos"], q["aos"]),
                          ("rgb", p["result_cloud_color"], o["rgb"], q["rgb"]), ("map", p["cloud_pose_map"], o["pose_map"], q["pose_map"]),
                          ("dc", p["dc_index"], o["dc_index"], q["dc_index"])]:
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), f"{name}: product != oracle"
        assert np.array_equal(b.view(np.uint8), c.view(np.uint8)), f"{name}: oracle != reference"
    if mode == "label":
        assert np.array_equal(p["result_observed_cloud_label"], o["label"])
        assert np.array_equal(o["label"], q["label"])


# ---------------------------------------------------------------------------------------------------------------------
def _cost_case(orc, n_tris, stride, dof6, n_poses, seed=42):
    case = make_case(n_tris, 3, seed)
    s = case.scene
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    poses, pmm, lab = poses_near_gt(case, n_poses)
    if dof6:
        cnt = np.bincount(obs["label"], minlength=len(s.gt_poses)).astype(np.float32)
        tot = cnt[lab]
    else:
        tot = synth.cylinder_counts(s, obs["aos"], poses, pmm)
    return case, obs, poses, pmm, (lab if dof6 else None), tot


@pytest.mark.parametrize("n_tris,stride,dof6", [(600, 4, False), (5000, 4, False), (5000, 8, True), (600, 5, False), (2000, 1, True)])
def test_cost_parity(orc, ref, n_tris, stride, dof6):
    n_poses = 12 if stride == 1 else 40
    case, obs, poses, pmm, lab, tot = _cost_case(orc, n_tris, stride, dof6, n_poses)
    s = case.scene
    res = 0.01 if dof6 else 0.0075
    ct = 2 if dof6 else 0
    with make_renderer(case, with_label=dof6) as r:
        r.set_observed(obs["aos"], obs["rgb"], obs["label"] if dof6 else None)
        out = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, stride=stride, cost_type=ct, sensor_resolution=res,
                                          occlusion_threshold=1.0)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          s.source_label if dof6 else None, lab, 1.0)
    o = orc.costs(d_o, c_o, case.width, case.height, case.cx, case.cy, case.fx, case.fy, 100.0, stride, lab, obs["aos"],
                  obs["rgb"], obs["label"] if dof6 else None, tot, ct, True, res, 20.0)
    assert np.array_equal(out["counts"], o["counts"]), f"counts differ:\n{out['counts']}\n{o['counts']}"
    assert o["counts"][0].sum() > 0 and o["counts"][2].sum() > 0
    for k in ("rendered_cost", "observed_cost", "points_diff_cost"):
        assert np.array_equal(out[k], o[k], equal_nan=True), k
    # the reference's own compute_point_clouds + compute_costs kernels on the canonical images: exact
    q = ref.cost_from_images(d_o, c_o, case.width, case.height, case.cx, case.cy, case.fx, case.fy, 100, stride, lab,
                             obs["aos"], obs["rgb"], obs["label"] if dof6 else None, tot, ct, True, res, 20.0)
    assert q["n_rend"] == int(o["counts"][0].sum())
    for k in ("rendered_cost", "observed_cost", "points_diff_cost"):
        np.testing.assert_allclose(out[k], q[k], rtol=1e-4, atol=0, equal_nan=True, err_msg=f"{k} vs reference kernels")
        assert np.array_equal(out[k], q[k], equal_nan=True), k
    # the reference end to end (its racy rasterizer included): reported, loosely bounded
    e = ref.render_cost(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                        s.source_color, s.source_label if dof6 else None, lab, stride, 100, case.cx, case.cy, case.fx,
                        case.fy, obs["aos"], obs["rgb"], obs["label"] if dof6 else None, tot, ct, True, res, 20.0, 1.0)
    dev = np.abs(e["rendered_cost"] - out["rendered_cost"])
    print(f"reference end-to-end rendered_cost deviation from canonical: median {np.median(dev):.3f}, max {dev.max():.3f}")
    assert np.median(dev) < 25.0  # the deviation is the reference's own nondeterminism, not ours


def test_cost_colour_flip_tolerant(orc, ref):
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 5000, 4, False, 32)
    s = case.scene
    thr = 20.0
    with make_renderer(case) as r:
        r.set_observed(obs["aos"], obs["rgb"], None)
        out = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, cost_type=1, sensor_resolution=0.0075,
                                          color_distance_threshold=thr)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          None, None, 1.0)
    o = orc.costs(d_o, c_o, case.width, case.height, case.cx, case.cy, case.fx, case.fy, 100.0, 4, None, obs["aos"],
                  obs["rgb"], None, tot, 1, True, 0.0075, thr, color_eps=1e-3)
    q = ref.cost_from_images(d_o, c_o, case.width, case.height, case.cx, case.cy, case.fx, case.fy, 100, 4, None, obs["aos"],
                             obs["rgb"], None, tot, 1, True, 0.0075, thr)
    assert np.array_equal(out["counts"][0], o["counts"][0])
    # rendered points whose CIEDE2000 distance is within 1e-3 of the threshold may flip (float libm differs CPU/GPU)
    slack = o["near_threshold"]
    assert np.all(np.abs(out["counts"][1] - o["counts"][1]) <= slack)
    assert np.all(np.abs(out["counts"][2] - o["counts"][2]) <= slack)
    assert (o["counts"][1] > 0).any() and (o["counts"][2] > 0).any()
    # colour matters: depth-only scoring of the same poses explains more points
    assert (out["counts"][1] >= _depth_only_counts(case, obs, poses, pmm, tot)[1]).all()
    # same GPU, same math library: equal to the reference kernels
    np.testing.assert_allclose(out["rendered_cost"], q["rendered_cost"], rtol=1e-4)
    np.testing.assert_allclose(out["observed_cost"], q["observed_cost"], rtol=1e-4)


def _depth_only_counts(case, obs, poses, pmm, tot):
    with make_renderer(case) as r:
        r.set_observed(obs["aos"], obs["rgb"], None)
        return r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, cost_type=0, sensor_resolution=0.0075)["counts"]


def test_cloud_stage_parity(orc, ref):
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 5000, 4, False, 20)
    s = case.scene
    with make_renderer(case) as r:
        out = r.render_cuda_multi_unified("CLOUD", poses, pmm, stride=4)
    d_o, c_o = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          None, None, 1.0)
    o = orc.depth2cloud(d_o, c_o, case.width, case.height, len(poses), case.cx, case.cy, case.fx, case.fy, 100.0, 4)
    assert out["result_cloud_point_num"] == o["n"] > 0
    assert np.array_equal(out["result_cloud"].view(np.uint32), o["soa"].view(np.uint32))
    assert np.array_equal(out["result_cloud_color"], o["rgb"])
    assert np.array_equal(out["result_cloud_pose_map"], o["pose_map"])
    q = ref.depth2cloud(d_o, c_o, case.width, case.height, len(poses), case.cx, case.cy, case.fx, case.fy, 100.0, 4)
    assert q["n"] == o["n"]
    assert np.array_equal(q["soa"].view(np.uint32), o["soa"].view(np.uint32))
    assert np.array_equal(q["pose_map"], o["pose_map"])


def test_debug_stage_returns_everything(orc):
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 600, 4, False, 6)
    with make_renderer(case) as r:
        r.set_observed(obs["aos"], obs["rgb"], None)
        dbg = r.render_cuda_multi_unified("DEBUG", poses, pmm, None, tot, stride=4, sensor_resolution=0.0075)
        cost = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, sensor_resolution=0.0075)
        img = r.render_cuda_multi_unified("RENDER", poses, pmm)
    assert np.array_equal(dbg["result_depth"], img["result_depth"])
    assert np.array_equal(dbg["counts"], cost["counts"])
    assert dbg["result_cloud_point_num"] == int(cost["counts"][0].sum())


def test_cost_edge_cases(orc):
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 600, 4, False, 8)
    s = case.scene
    far = poses[0].copy(); far[0, 3] += 500.0  # 5 m to the right: nothing rendered
    poses2 = np.concatenate([poses, far[None]])
    pmm2 = np.concatenate([pmm, pmm[:1]])
    tot2 = np.concatenate([tot, [0.0]]).astype(np.float32)
    with make_renderer(case) as r:
        r.set_observed(obs["aos"], obs["rgb"], None)
        out = r.render_cuda_multi_unified("COST", poses2, pmm2, None, tot2, stride=4, sensor_resolution=0.0075)
        assert out["rendered_cost"][-1] == -1.0  # compute_costs.cuh:28-31
        assert np.isnan(out["observed_cost"][-1])  # 0/0, compute_costs.cuh:435-439
        assert out["counts"][0, -1] == 0
        empty = r.render_cuda_multi_unified("COST", np.zeros((0, 4, 4), np.float32), np.zeros(0, np.int32), None,
                                            np.zeros(0, np.float32), stride=4)
        assert empty["rendered_cost"].shape == (0,)
        # batch-split invariance: poses are independent
        a = r.render_cuda_multi_unified("COST", poses2[:4], pmm2[:4], None, tot2[:4], stride=4, sensor_resolution=0.0075)
        b = r.render_cuda_multi_unified("COST", poses2[4:], pmm2[4:], None, tot2[4:], stride=4, sensor_resolution=0.0075)
        assert np.array_equal(np.concatenate([a["counts"], b["counts"]], 1), out["counts"])
        # empty observed cloud: every rendered point unexplained
        r.set_observed(np.zeros((0, 3), np.float32), None, None)
        none = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, sensor_resolution=0.0075)
        assert np.array_equal(none["counts"][0], none["counts"][1]) and none["counts"][2].sum() == 0
        with pytest.raises(Exception):
            r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=7)  # 640 % 7 != 0


def test_forced_bands_and_tiny_queues(orc, monkeypatch):
    """A tile cap far below the tile size cuts every pose into many bands (and IMAGE mode likewise): results must not
    change.  Exercises the band clamp of the culling tests and the multi-round candidate queue."""
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 5000, 4, False, 24)
    with make_renderer(case) as r:
        r.set_observed(obs["aos"], obs["rgb"], None)
        a = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, sensor_resolution=0.0075)
        ia = r.render_cuda_multi_unified("RENDER", poses[:6], pmm[:6])
        monkeypatch.setenv("PERCH_COST_TILE_CAP", "1")   # clamped up to one lattice row per band
        monkeypatch.setenv("PERCH_IMAGE_TILE_CAP", "1")
        b = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, sensor_resolution=0.0075)
        ib = r.render_cuda_multi_unified("RENDER", poses[:6], pmm[:6])
    assert a["counts"][0].sum() > 0
    assert np.array_equal(a["counts"], b["counts"])
    assert np.array_equal(ia["result_depth"], ib["result_depth"]) and np.array_equal(ia["result_color"], ib["result_color"])


def test_non_finite_vertices_and_degenerate_triangles(orc):
    """NaN / inf vertices and zero-area triangles must neither crash nor leak through the conservative culling: the
    exact path decides, exactly like the reference's float sequence (oracle)."""
    case = make_case(600, 3, 42)
    s = case.scene
    tris = s.tris.copy()
    n0 = int(s.tris_model_count[0])
    rng = np.random.default_rng(5)
    bad = rng.choice(n0, 24, replace=False)
    v = tris["v"]
    v[bad[:6], 0, 0] = np.nan
    v[bad[6:12], 1, 2] = np.inf
    v[bad[12:18], 2] = v[bad[12:18], 1]          # two equal vertices: zero area
    v[bad[18:24]] = v[bad[18:24], :1]            # a point
    poses, pmm, _ = poses_near_gt(case, 8)
    with make_renderer(case) as r:
        r.set_models(tris, s.tris_model_count)
        out = r.render_cuda_multi_unified("RENDER", poses, pmm)
    d_o, c_o = orc.render(tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                          None, None, 1.0)
    assert (d_o > 0).sum() > 0
    assert np.array_equal(out["result_depth"], d_o)
    assert np.array_equal(out["result_color"], c_o)


@pytest.mark.parametrize("stage", ["RENDER", "COST"])
def test_edge_grazing_triangles(orc, stage):
    """Adversarial input for the culling levels (pose_kernels.cu pass A / level 3): small triangles whose edges pass
    within 0 ... 0.05 px of the one pixel their bounding box holds, near-degenerate slivers, and triangles with a vertex
    on a pixel centre.  Whether such a fragment exists is decided by the reference's float sequence alone (oracle): the
    approximate tests may only ever drop triangles that sequence rejects."""
    from perception_b200.renderer import make_triangles, pose_to_mat4x4
    case = make_case(600, 3, 42)
    W, H, P = case.width, case.height, case.proj
    rng = np.random.default_rng(11)
    step = 1 if stage == "RENDER" else 4
    n = 6000
    px = rng.integers(8, W // step - 8, n) * step
    py = (H - 1) - rng.integers(8, H // step - 8, n) * step           # raster row of a lattice image row (image row = H-1-P1)
    Z = rng.uniform(40.0, 90.0, n)                                    # cm
    eps = rng.choice([0.0, 1e-5, 1e-4, 1e-3, 1e-2, 0.05, 0.3], n) * rng.choice([-1.0, 1.0], n)
    ang = rng.uniform(0, 2 * np.pi, n)
    size = rng.uniform(0.3, 0.9, n)
    sliver = rng.uniform(size=n) < 0.2
    # an edge through (px, py) + eps * normal, direction `ang`; third vertex on either side
    d = np.stack([np.cos(ang), np.sin(ang)], 1)
    nrm = np.stack([-d[:, 1], d[:, 0]], 1)
    c = np.stack([px, py], 1).astype(np.float64) + eps[:, None] * nrm
    a = c - d * size[:, None] * rng.uniform(0.2, 1.0, n)[:, None]
    b = c + d * size[:, None] * rng.uniform(0.2, 1.0, n)[:, None]
    third = c + nrm * np.where(sliver, 1e-3, size)[:, None] * rng.choice([-1.0, 1.0], n)[:, None] + d * rng.uniform(-0.3, 0.3, n)[:, None]
    on_vertex = rng.uniform(size=n) < 0.1
    a[on_vertex] = np.stack([px, py], 1)[on_vertex]
    scr = np.stack([a, b, third], 1)                                  # (n, 3, 2) screen coordinates (raster rows)
    # un-project: sx = (P.a0*X + P.a2*Z)/Z * W/2 + W/2, sy likewise with row b (image_renderer.cuh:82-84)
    X = ((scr[..., 0] - W / 2) / (W / 2) - P[0, 2]) * Z[:, None] / P[0, 0]
    Y = ((scr[..., 1] - H / 2) / (H / 2) - P[1, 2]) * Z[:, None] / P[1, 1]
    verts = np.stack([X, Y, np.broadcast_to(Z[:, None], X.shape)], -1) / 100.0   # metres; pose scales by 100
    tris = make_triangles(verts.astype(np.float32), rng.integers(0, 255, (n, 3)))
    counts = np.array([n], np.int32)
    poses = pose_to_mat4x4(np.eye(4))[None]
    pmm = np.zeros(1, np.int32)
    src = np.zeros((H, W), np.int32)
    with make_renderer(case) as r:
        r.set_models(tris, counts)
        r.set_scene(src)
        if stage == "RENDER":
            out = r.render_cuda_multi_unified("RENDER", poses, pmm)
            d_o, c_o = orc.render(tris, counts, poses, pmm, W, H, P, src, None, None, 1.0)
            assert 500 < (d_o > 0).sum() < n
            assert np.array_equal(out["result_depth"], d_o)
            assert np.array_equal(out["result_color"], c_o)
        else:
            out = r.render_cuda_multi_unified("CLOUD", poses, pmm, stride=step)
            d_o, c_o = orc.render(tris, counts, poses, pmm, W, H, P, src, None, None, 1.0)
            want = orc.depth2cloud(d_o[0], c_o[:, 0], W, H, 1, case.cx, case.cy, case.fx, case.fy, 100.0, step)
            assert want["n"] > 300
            assert out["result_cloud_point_num"] == want["n"]
            assert np.array_equal(out["result_cloud"], want["soa"])
            assert np.array_equal(out["result_cloud_color"], want["rgb"])


def test_pose_permutation_and_duplicates(orc):
    """Poses are independent units: permuting the pose list permutes the results, a duplicated pose scores the same."""
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 2000, 8, True, 32)
    perm = np.random.default_rng(1).permutation(len(poses))
    poses2 = np.concatenate([poses[perm], poses[:5]])
    pmm2 = np.concatenate([pmm[perm], pmm[:5]])
    lab2 = np.concatenate([lab[perm], lab[:5]])
    tot2 = np.concatenate([tot[perm], tot[:5]])
    with make_renderer(case, with_label=True) as r:
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])
        a = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, stride=8, cost_type=2, sensor_resolution=0.01)
        b = r.render_cuda_multi_unified("COST", poses2, pmm2, lab2, tot2, stride=8, cost_type=2, sensor_resolution=0.01)
    n = len(poses)
    assert np.array_equal(b["counts"][:, :n], a["counts"][:, perm])
    assert np.array_equal(b["counts"][:, n:], a["counts"][:, :5])
    assert np.array_equal(b["rendered_cost"][:n], a["rendered_cost"][perm])


def test_triangle_order_invariance():
    """Depth is a min over fragments: shuffling the triangle soup must not change it (colour may, on ties)."""
    case = make_case(5000, 3, 42)
    s = case.scene
    poses, pmm, _ = poses_near_gt(case, 8)
    keep = pmm == 0
    poses, pmm = poses[keep], pmm[keep]
    n0 = int(s.tris_model_count[0])
    perm = np.random.default_rng(0).permutation(n0)
    tris2 = s.tris.copy()
    tris2[:n0] = s.tris[:n0][perm]
    with make_renderer(case) as r:
        a = r.render_cuda_multi_unified("RENDER", poses, pmm)["result_depth"]
        r.set_models(tris2, s.tris_model_count)
        b = r.render_cuda_multi_unified("RENDER", poses, pmm)["result_depth"]
    assert np.array_equal(a, b)


def test_multi_model_blocks_c5_shape(orc):
    """BASELINE configs[4] in shape (many meshes, pose list in per-model blocks, labels): every model's block must score
    exactly as it does in a context that holds that mesh alone, and a sample must equal the oracle."""
    M, T, per = 7, 1500, 48
    meshes = [synth.make_mesh(T, seed=1000 + m, r0=0.05 + 0.01 * m) for m in range(M)]
    K = synth.camera_K()
    proj = orc.compute_proj(K, 640, 480)

    def render_fn(tris, counts, poses, pmm):
        with Renderer(0) as rr:
            rr.set_camera(640, 480, proj, synth.CX, synth.CY, synth.FX, synth.FY, 100.0)
            rr.set_models(tris, counts)
            rr.set_scene(np.zeros(640 * 480, np.int32))
            o = rr.render_cuda_multi_unified("RENDER", poses, pmm)
        return o["result_depth"], o["result_color"]

    scene = synth.make_scene(meshes, render_fn, n_instances=M, seed=5)
    P = M * per
    pmm = np.repeat(np.arange(M, dtype=np.int32), per)
    poses = np.zeros((P, 4, 4), np.float32)
    lab = np.zeros(P, np.int32)
    for m in range(M):
        # a model the scene has no room for borrows another instance's place and label
        g = scene.gt_models.index(m) if m in scene.gt_models else m % len(scene.gt_poses)
        poses[pmm == m] = synth.poses_6dof(scene.gt_poses[g][:3, 3], per, n_views=12, n_inplane=4, seed=m)
        poses[m * per] = pose_to_mat4x4(scene.gt_poses[g])
        lab[pmm == m] = g
    kw = dict(stride=8, cost_type=2, sensor_resolution=0.01)
    cam = (synth.CX, synth.CY, synth.FX, synth.FY, 100.0)
    with Renderer(0) as r:
        r.set_camera(640, 480, proj, *cam)
        r.set_models(scene.tris, scene.tris_model_count)
        r.set_scene(scene.source_depth, scene.source_color, scene.source_label)
        obs = r.depth2cloud_global(scene.source_depth, scene.source_color, 640, 480, 1, *cam, 8, scene.source_label)
        olab = obs["result_observed_cloud_label"]
        r.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], olab)
        tot = np.bincount(olab, minlength=len(scene.gt_poses)).astype(np.float32)[lab]
        full = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        sel = np.arange(0, P, 29)
        img = r.render_cuda_multi_unified("RENDER", poses[sel], pmm[sel], lab[sel])
    assert len(scene.gt_poses) >= 4 and (full["counts"][0].reshape(M, per)[:, 0] > 20).all()  # first pose of a block = an instance's pose
    o = orc.costs(img["result_depth"], img["result_color"], 640, 480, *cam, 8, lab[sel], obs["result_cloud_eigen"],
                  obs["result_cloud_color"], olab, tot[sel], 2, True, 0.01, 20.0)
    assert np.array_equal(full["counts"][:, sel], o["counts"])
    for m in range(M):
        blk = pmm == m
        with Renderer(0) as r1:
            r1.set_camera(640, 480, proj, *cam)
            r1.set_models(meshes[m], [T])
            r1.set_scene(scene.source_depth, scene.source_color, scene.source_label)
            r1.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], olab)
            one = r1.render_cuda_multi_unified("COST", poses[blk], np.zeros(per, np.int32), lab[blk], tot[blk], **kw)
        assert np.array_equal(one["counts"], full["counts"][:, blk]), m
        assert np.array_equal(one["rendered_cost"], full["rendered_cost"][blk])


def test_full_size_properties(orc):
    """BASELINE config-2 shape at full size (20k-triangle mesh, 640x480, stride 8, labels): the oracle is too slow
    here, so check size-independent properties: COST counts == counts recomputed from the RENDER stage images by the
    oracle's cloud+cost code on a sample, determinism, and top-k consistency."""
    mesh = synth.make_mesh(20000, seed=1000, r0=0.09)
    case = make_case(600, 3, 42)  # camera only
    K = synth.camera_K()
    proj = orc.compute_proj(K, 640, 480)

    def render_fn(tris, counts, poses, pmm):
        with Renderer(0) as rr:
            rr.set_camera(640, 480, proj, synth.CX, synth.CY, synth.FX, synth.FY, 100.0)
            rr.set_models(tris, counts)
            rr.set_scene(np.zeros(640 * 480, np.int32))
            o = rr.render_cuda_multi_unified("RENDER", poses, pmm)
        return o["result_depth"], o["result_color"]

    scene = synth.make_scene([mesh], render_fn, n_instances=3, seed=11)
    P = 2048
    center = scene.gt_poses[0][:3, 3]
    poses = synth.poses_6dof(center, P, n_views=64, n_inplane=32)
    poses[0] = pose_to_mat4x4(scene.gt_poses[0])
    pmm = np.zeros(P, np.int32)
    lab = np.zeros(P, np.int32)
    with Renderer(0) as r:
        r.set_camera(640, 480, proj, synth.CX, synth.CY, synth.FX, synth.FY, 100.0)
        r.set_models(scene.tris, scene.tris_model_count)
        r.set_scene(scene.source_depth, scene.source_color, scene.source_label)
        obs = r.depth2cloud_global(scene.source_depth, scene.source_color, 640, 480, 1, synth.CX, synth.CY, synth.FX,
                                   synth.FY, 100.0, 8, scene.source_label)
        r.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], obs["result_observed_cloud_label"])
        tot = np.full(P, float((obs["result_observed_cloud_label"] == 0).sum()), np.float32)
        a = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, stride=8, cost_type=2, sensor_resolution=0.01)
        b = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, stride=8, cost_type=2, sensor_resolution=0.01)
        assert np.array_equal(a["counts"], b["counts"])  # deterministic
        sel = np.arange(0, P, 128)
        img = r.render_cuda_multi_unified("RENDER", poses[sel], pmm[sel], lab[sel])
    o = orc.costs(img["result_depth"], img["result_color"], 640, 480, synth.CX, synth.CY, synth.FX, synth.FY, 100.0, 8,
                  lab[sel], obs["result_cloud_eigen"], obs["result_cloud_color"], obs["result_observed_cloud_label"], tot[sel],
                  2, True, 0.01, 20.0)
    assert np.array_equal(a["counts"][:, sel], o["counts"])
    assert np.array_equal(a["rendered_cost"][sel], o["rendered_cost"])
    # the ground-truth pose must be the best-scoring candidate region
    best = topk(a["rendered_cost"], a["observed_cost"], pmm, 1, k=5)
    assert best["cost"][0, 0] <= int(a["rendered_cost"][0] + a["observed_cost"][0])
    assert a["rendered_cost"][0] < 35.0


def test_observed_order_within_a_label_does_not_change_counts(orc):
    """The reference sorts the observed cloud by label with thrust::sort_by_key (renderer.cu:1681), which leaves the order
    INSIDE a label unspecified; this library and the oracle harness sort stably.  Nearest-neighbour indices (and their
    tie-breaks) follow that order, but `explained` is a set of POINTS, so for cost_type 0 / 2 every count and cost must
    be the same for any order -- checked by shuffling the observed cloud (with exact duplicates in it, i.e. real distance
    ties) before handing it over.  (cost_type 1 looks colours up by index in the UNSORTED colour array, renderer.cu:1881:
    there the order matters in the reference itself, which is why it is not part of this property.)"""
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 5000, 8, True, 40)
    rng = np.random.default_rng(11)
    xyz, rgb, olab = obs["aos"].copy(), obs["rgb"].copy(), obs["label"].copy()
    dup = rng.choice(len(xyz), 60, replace=False)          # exact duplicates => equal distances => index tie-breaks
    xyz = np.concatenate([xyz, xyz[dup]]); rgb = np.concatenate([rgb, rgb[:, dup]], 1); olab = np.concatenate([olab, olab[dup]])
    tot = np.bincount(olab, minlength=lab.max() + 1).astype(np.float32)[lab]
    outs = []
    for trial in range(3):
        perm = np.arange(len(xyz)) if trial == 0 else rng.permutation(len(xyz))
        with make_renderer(case, with_label=True) as r:
            r.set_observed(xyz[perm], rgb[:, perm], olab[perm])
            outs.append(r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, stride=8, cost_type=2, sensor_resolution=0.01))
    assert outs[0]["counts"][2].sum() > 0
    for o in outs[1:]:
        assert np.array_equal(o["counts"], outs[0]["counts"])
        for k in ("rendered_cost", "observed_cost", "points_diff_cost"):
            assert np.array_equal(o[k], outs[0][k], equal_nan=True), k


@pytest.mark.gpu
def test_set_models_is_idempotent_and_notices_changes(orc):
    """The reference passes the meshes with every render call (renderer.cu:1532-1544), so the drop-in backend calls
    perch_set_models every time: an unchanged mesh set must be recognised (digest) and left in place, a changed one of
    the same size must be re-filed -- one moved vertex is enough to change the result."""
    import time
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 5000, 8, True, 24)
    s = case.scene
    kw = dict(stride=8, cost_type=2, sensor_resolution=0.01)
    moved = s.tris.copy()
    big = int(np.argmax(s.tris_model_count))
    lo = int(np.sum(s.tris_model_count[:big]))
    moved["v"][lo:lo + int(s.tris_model_count[big]), 0, :] += np.float32(0.02)   # first vertex of every triangle of one model: 2 cm
    with make_renderer(case, with_label=True) as r:
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])
        a = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        t0 = time.perf_counter(); r.set_models(s.tris, s.tris_model_count); t_same = time.perf_counter() - t0
        a2 = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        t0 = time.perf_counter(); r.set_models(moved, s.tris_model_count); t_new = time.perf_counter() - t0
        b = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
    with make_renderer(case, with_label=True) as r:
        r.set_models(moved, s.tris_model_count)
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])
        b_fresh = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
    assert np.array_equal(a["counts"], a2["counts"])
    assert np.array_equal(b["counts"], b_fresh["counts"]) and np.array_equal(b["rendered_cost"], b_fresh["rendered_cost"])
    assert not np.array_equal(a["counts"], b["counts"]), "the moved mesh must score differently"
    assert t_same < t_new, (t_same, t_new)


@pytest.mark.gpu
def test_set_observed_is_idempotent_and_notices_changes(orc):
    """Same contract for the observed cloud (passed with every call by the reference, renderer.cu:1668): an identical
    cloud keeps the resident copy and its grids, a cloud that differs in one coordinate, colour or label replaces them."""
    case, obs, poses, pmm, lab, tot = _cost_case(orc, 5000, 8, True, 24)
    kw = dict(stride=8, cost_type=2, sensor_resolution=0.01)
    moved = obs["aos"].copy()
    moved[: len(moved) // 2, 2] += np.float32(0.05)          # half the cloud 5 cm farther away
    with make_renderer(case, with_label=True) as r:
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])
        a = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])  # unchanged: no-op
        a2 = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        r.set_observed(moved, obs["rgb"], obs["label"])
        b = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])  # and back
        a3 = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
    with make_renderer(case, with_label=True) as r:
        r.set_observed(moved, obs["rgb"], obs["label"])
        b_fresh = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
    for x in (a2, a3):
        assert np.array_equal(a["counts"], x["counts"]) and np.array_equal(a["observed_cost"], x["observed_cost"], equal_nan=True)
    assert np.array_equal(b["counts"], b_fresh["counts"])
    assert not np.array_equal(a["counts"], b["counts"]), "the moved cloud must score differently"
