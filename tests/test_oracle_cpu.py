"""CPU tests of the oracle (test infrastructure): known-answer cases derived by hand from the reference's formulas,
structural properties, and the committed golden vectors produced by the reference's own CUDA kernels
(tests/golden/make_golden.py, run on a B200)."""
from pathlib import Path

import numpy as np
import pytest

from oracle_lib import Oracle
from perception_b200 import synth
from perception_b200.renderer import make_triangles, pose_to_mat4x4
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt

GOLDEN = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def test_compute_proj_known_answer(orc):
    # renderer.cu:1386-1410 with the kinect intrinsics of object_recognizer.cpp:64-75
    P = orc.compute_proj(synth.camera_K(), 640, 480, 10, 10000)
    f32 = np.float32
    assert P[0, 0] == f32(2) * f32(synth.FX) / f32(640)
    assert P[1, 1] == -(f32(2) * f32(synth.FY) / f32(480))
    assert P[0, 2] == -(f32(-2) * f32(synth.CX) / f32(640) + f32(1))
    assert P[1, 2] == -(f32(2) * f32(synth.CY) / f32(480) - f32(1))
    assert P[2, 2] == f32(10010) / f32(9990) and P[3, 2] == 1 and P[3, 3] == 0
    assert P[0, 1] == 0 and P[0, 3] == 0 and P[1, 0] == 0


def _quad(z_m, half=0.05, color=(10, 20, 30)):
    v = np.array([[[-half, -half, z_m], [half, -half, z_m], [half, half, z_m]],
                  [[-half, -half, z_m], [half, half, z_m], [-half, half, z_m]]], np.float32)
    return make_triangles(v, np.array([color, color]))


def test_render_fronto_parallel_quad(orc):
    """A 10 cm square at z = 0.80 m rendered with the identity pose (x100): every covered pixel has depth 80 cm, its
    extent matches the pinhole projection with the reference's one-row vertical offset (row = v - 1, because
    y_to_write = H-1-P1 while sy = H - v; image_renderer.cuh:82-84,128)."""
    W, H = 640, 480
    proj = orc.compute_proj(synth.camera_K(), W, H)
    tris = _quad(0.80)
    d, rgb = orc.render(tris, [2], pose_to_mat4x4(np.eye(4))[None], [0], W, H, proj)
    ys, xs = np.nonzero(d[0])
    assert set(np.unique(d[0])) == {0, 80}
    u0, u1 = synth.CX - synth.FX * 0.05 / 0.8, synth.CX + synth.FX * 0.05 / 0.8
    v0, v1 = synth.CY - synth.FY * 0.05 / 0.8, synth.CY + synth.FY * 0.05 / 0.8
    assert xs.min() == int(np.ceil(u0)) and xs.max() == int(np.floor(u1))
    assert ys.min() == int(np.ceil(v0)) - 1 and ys.max() == int(np.floor(v1)) - 1
    assert (rgb[0][0][d[0] > 0] == 10).all() and (rgb[1][0][d[0] > 0] == 20).all() and (rgb[2][0][d[0] > 0] == 30).all()
    assert (rgb[:, 0][:, d[0] == 0] == 0).all()


def test_render_ztest_and_lowest_index_colour(orc):
    W, H = 160, 120
    K = np.array([[144.0, 0, 80], [0, 144.0, 60], [0, 0, 1]], np.float32)
    proj = orc.compute_proj(K, W, H)
    near, far, tie = _quad(0.50, 0.03, (1, 1, 1)), _quad(0.70, 0.06, (2, 2, 2)), _quad(0.50, 0.03, (3, 3, 3))
    tris = np.concatenate([far, near, tie])
    d, rgb = orc.render(tris, [6], pose_to_mat4x4(np.eye(4))[None], [0], W, H, proj)
    assert set(np.unique(d[0])) == {0, 50, 70}
    assert (rgb[0][0][d[0] == 50] == 1).all()  # `near` (index 2,3) beats `tie` (index 4,5) at equal depth: strict <
    assert (rgb[0][0][d[0] == 70] == 2).all()
    # triangles of another model are skipped (image_renderer.cuh:242-243)
    d2, _ = orc.render(tris, [2, 4], pose_to_mat4x4(np.eye(4))[None], [0], W, H, proj)
    assert set(np.unique(d2[0])) == {0, 70}


def test_source_occlusion_rules(orc):
    """image_renderer.cuh:163-186: 3-DoF |d-s| > thr && d > s > 0 clears the pixel; 6-DoF needs a label mismatch."""
    W, H = 160, 120
    K = np.array([[144.0, 0, 80], [0, 144.0, 60], [0, 0, 1]], np.float32)
    proj = orc.compute_proj(K, W, H)
    tris = _quad(0.60, 0.05)
    pose = pose_to_mat4x4(np.eye(4))[None]
    base, _ = orc.render(tris, [2], pose, [0], W, H, proj)
    cov = base[0] > 0
    for s_val, thr, cleared in [(58, 1.0, True), (59, 1.0, False), (59, 0.5, True), (60, 0.0, False), (61, 0.0, False), (0, 0.0, False)]:
        src = np.full((H, W), s_val, np.int32)
        d, rgb = orc.render(tris, [2], pose, [0], W, H, proj, src, None, None, thr)
        assert ((d[0][cov] == 0).all() if cleared else (d[0][cov] == 60).all()), (s_val, thr)
        if cleared:
            assert (rgb[:, 0][:, cov] == 0).all()
    src = np.full((H, W), 55, np.int32)
    lab = np.zeros((H, W), np.uint8)
    lab[:, : W // 2] = 1  # label 1 == pose label 0 (source labels start at 1, :162)
    lab[:, W // 2:] = 2
    d, _ = orc.render(tris, [2], pose, [0], W, H, proj, src, lab, [0], 1.0)
    left = cov.copy(); left[:, W // 2:] = False
    right = cov.copy(); right[:, : W // 2] = False
    assert (d[0][left] == 60).all() and (d[0][right] == 0).all()


def test_render_order_independent(orc):
    case = make_case(600, 3, 42, 160, 120)
    s = case.scene
    poses, pmm, _ = poses_near_gt(case, 6)
    keep = pmm == 0
    n0 = int(s.tris_model_count[0])
    a, _ = orc.render(s.tris, s.tris_model_count, poses[keep], pmm[keep], 160, 120, case.proj, s.source_depth)
    t2 = s.tris.copy()
    t2[:n0] = s.tris[:n0][::-1]
    b, _ = orc.render(t2, s.tris_model_count, poses[keep], pmm[keep], 160, 120, case.proj, s.source_depth)
    assert np.array_equal(a, b) and (a > 0).sum() > 100


def test_depth2cloud_known_answer(orc):
    W, H = 16, 8
    depth = np.zeros((H, W), np.int32)
    depth[0, 0], depth[4, 8], depth[4, 12], depth[3, 3], depth[0, 4] = 100, 250, -5, 77, 50
    col = np.stack([np.full((H, W), k, np.uint8) for k in (7, 8, 9)])
    o = orc.depth2cloud(depth, col, W, H, 1, 8.0, 4.0, 10.0, 20.0, 100.0, 4, want_dc_index=True)
    assert o["n"] == 3  # (0,0), (4,0) on row 0, (8,4) on row 4; (3,3) is off-lattice, -5 is invalid
    f = np.float32
    exp = [(f(0 - 8) / f(10) * f(1.0), f(0 - 4) / f(20) * f(1.0), f(1.0)), (f(4 - 8) / f(10) * f(0.5), f(0 - 4) / f(20) * f(0.5), f(0.5)),
           (f(0), f(0), f(2.5))]
    np.testing.assert_array_equal(o["aos"], np.array(exp, np.float32))
    np.testing.assert_array_equal(o["soa"], o["aos"].T)
    assert (o["rgb"] == np.array([[7], [8], [9]])).all()
    dc = o["dc_index"].reshape(H, W)
    assert dc[0, 0] == 0 and dc[0, 1] == 1 and dc[0, 4] == 1 and dc[0, 5] == 2 and dc[4, 8] == 2 and dc[4, 9] == 3 and dc[7, 15] == 3
    lab = np.zeros((H, W), np.uint8)
    lab[0, 4], lab[4, 8] = 3, 1
    o2 = orc.depth2cloud(depth, col, W, H, 1, 8.0, 4.0, 10.0, 20.0, 100.0, 4, label_mask=lab)
    assert o2["n"] == 2 and list(o2["label"]) == [2, 0]


def test_depth2cloud_bounds(orc):
    case = make_case(600, 3, 42, 160, 120)
    o_all = orc.depth2cloud(case.scene.source_depth, None, 160, 120, 1, case.cx, case.cy, case.fx, case.fy, 100.0, 2)
    o_b = observed_3dof(case, 2, orc)
    assert 0 < o_b["n"] < o_all["n"]
    T_tc = np.linalg.inv(case.scene.T_cam_table)
    w = o_b["aos"].astype(np.float64) @ T_tc[:3, :3].T + T_tc[:3, 3]
    b = synth.table_bounds(case.scene)
    assert (w[:, 2] > b[5] - 1e-4).all() and (w[:, 2] < b[4] + 1e-4).all()


def test_knn_tie_break_and_label_ranges(orc):
    tgt = np.array([[0, 0, 0], [2, 0, 0], [2, 0, 0], [10, 0, 0]], np.float32)
    src = np.array([[1, 0, 0], [2.2, 0, 0], [100, 0, 0]], np.float32)
    d, i = orc.knn_k1(src, tgt)
    assert list(i) == [0, 1, 3] and d[0] == 1.0  # equidistant 0 and 1/2 -> lowest index; duplicate points -> first
    d, i = orc.knn_k1(src, tgt, src_label=[1, 0, 5], label_begin=[0, 2], label_end=[2, 4])
    assert list(i) == [2, 1, -1] and d[2] == np.finfo(np.float32).max


def test_costs_known_answer(orc):
    """Two 'rendered' pixels, three observed points: compute_costs.cuh formulas by hand."""
    W, H = 8, 4
    depth = np.zeros((2, H, W), np.int32)
    depth[0, 0, 0], depth[0, 0, 4], depth[1, 0, 4] = 100, 100, 100
    col = np.zeros((3, 2, H, W), np.uint8)
    cx, cy, fx, fy = 4.0, 2.0, 100.0, 100.0
    # observed: one point exactly at pixel (0,0)'s back-projection, one 5 mm from pixel (4,0)'s, one far away
    p00 = np.array([(0 - cx) / fx * 1.0, (0 - cy) / fy * 1.0, 1.0], np.float32)
    p40 = np.array([(4 - cx) / fx * 1.0, (0 - cy) / fy * 1.0, 1.0], np.float32)
    obs = np.stack([p00, p40 + np.array([0.005, 0, 0], np.float32), np.array([5, 5, 5], np.float32)])
    tot = np.array([3, 3], np.float32)
    r = orc.costs(depth, col, W, H, cx, cy, fx, fy, 100.0, 4, None, obs, None, None, tot, 0, True, 0.0075, 20.0)
    assert r["counts"].tolist() == [[2, 1], [0, 0], [2, 1]]
    np.testing.assert_array_equal(r["rendered_cost"], [0, 0])
    np.testing.assert_allclose(r["observed_cost"], [100 * (3 - 2) / 3, 100 * (3 - 1) / 3], rtol=1e-6)
    np.testing.assert_array_equal(r["points_diff_cost"], [0, 0])
    r = orc.costs(depth, col, W, H, cx, cy, fx, fy, 100.0, 4, None, obs, None, None, tot, 0, True, 0.004, 20.0)
    assert r["counts"].tolist() == [[2, 1], [1, 1], [1, 0]]
    np.testing.assert_array_equal(r["rendered_cost"], [50, 100])
    # a pose with no rendered point: -1 (compute_costs.cuh:28-31); zero observed total: NaN (0/0)
    depth[1] = 0
    r = orc.costs(depth, col, W, H, cx, cy, fx, fy, 100.0, 4, None, obs, None, None, np.array([3, 0], np.float32), 0, True, 0.0075, 20.0)
    assert r["rendered_cost"][1] == -1 and np.isnan(r["observed_cost"][1])


def test_colour_distance_matches_published_ciede2000(orc):
    """Sanity of the CIEDE2000 restatement on grey/colour pairs: symmetric-ish, zero on identical colours, and larger
    for visibly different colours (the reference's variant is float/double mixed, so only coarse checks)."""
    assert orc.color_distance_rgb((120, 80, 40), (120, 80, 40)) == 0
    a = orc.color_distance_rgb((120, 80, 40), (125, 82, 43))
    b = orc.color_distance_rgb((120, 80, 40), (20, 200, 90))
    assert 0 < a < 5 < b
    assert abs(orc.color_distance_rgb((255, 255, 255), (0, 0, 0)) - 100.0) < 1e-3


def test_scene_scores_rank_ground_truth_first(orc):
    case = make_case(600, 3, 42)
    s = case.scene
    obs = observed_3dof(case, 4, orc)
    poses, pmm, _ = poses_near_gt(case, 12)
    tot = synth.cylinder_counts(s, obs["aos"], poses, pmm)
    d, c = orc.render(s.tris, s.tris_model_count, poses, pmm, 640, 480, case.proj, s.source_depth)
    r = orc.costs(d, c, 640, 480, case.cx, case.cy, case.fx, case.fy, 100.0, 4, None, obs["aos"], obs["rgb"], None, tot, 0,
                  True, 0.0075, 20.0)
    n_gt = len(s.gt_poses)
    assert (r["rendered_cost"][:n_gt] < 5).all()
    assert r["rendered_cost"][:n_gt].max() < np.sort(r["rendered_cost"][n_gt:])[1]


import make_golden_inputs as gi  # noqa: E402


@pytest.mark.parametrize("name", gi.NAMES)
def test_oracle_matches_reference_golden(orc, name):
    """Golden vectors = outputs of the reference's own CUDA kernels (oracle/_ref) on seeded inputs, generated on a
    B200 by tests/golden/make_golden.py.  This is what pins the oracle.  EXACT cases must match bit for bit; the
    full-mesh renders are bounded by the race budget recorded at generation time (the reference rasterizer is not
    deterministic on sm_70+, see tests/make_golden_inputs.py)."""
    path = GOLDEN / f"{name}.npz"
    if not path.exists():
        pytest.skip(f"{path.name} not generated yet (run tests/golden/make_golden.py on a GPU box)")
    g = np.load(path)
    out = gi.run_oracle(orc, name)
    keys = [k for k in g.files if k.startswith("ref_")]
    assert keys
    for k in keys:
        a, b = g[k], out[k[4:]]
        assert a.shape == b.shape, k
        if name in gi.EXACT:
            assert np.array_equal(np.atleast_1d(a).view(np.uint8), np.atleast_1d(b).astype(a.dtype).view(np.uint8)) \
                or np.array_equal(a, b, equal_nan=True), \
                f"{k}: {np.count_nonzero(a != b)} of {a.size} values differ from the reference kernels"
        elif k == "ref_depth":
            bad = np.count_nonzero(a != b)
            assert bad <= int(g["race_budget"]), f"{bad} depth pixels differ, budget {int(g['race_budget'])}"
            # wherever they agree on depth the colour must agree too, except equal-depth ties
            same = a == b
            assert np.count_nonzero((g["ref_rgb"] != out["rgb"]).any(0) & same) <= int(g["race_budget"])


@pytest.mark.parametrize("dof6", [False, True])
def test_score_poses_equals_render_then_costs(dof6):
    """oracle_score_poses (per pose, OpenMP, no images kept -- what the full-size GPU parity tests and bench.py's
    parity_check call) == oracle_render + oracle_costs on the whole batch, colour cost included."""
    from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt
    orc = Oracle()
    case = make_case(600, 3, 42)
    s = case.scene
    stride = 8 if dof6 else 4
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    poses, pmm, lab = poses_near_gt(case, 12)
    lab = lab if dof6 else None
    slab = s.source_label if dof6 else None
    olab = obs["label"] if dof6 else None
    tot = np.full(12, 100, np.float32)
    ct = 2 if dof6 else 1
    d, c = orc.render(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth, slab, lab)
    a = orc.costs(d, c, case.width, case.height, case.cx, case.cy, case.fx, case.fy, 100.0, stride, lab, obs["aos"], obs["rgb"],
                  olab, tot, ct, True, 0.01, 20.0)
    b = orc.score_poses(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth, slab, lab,
                        case.cx, case.cy, case.fx, case.fy, 100.0, stride, obs["aos"], obs["rgb"], olab, tot, ct, True, 0.01, 20.0)
    assert a["counts"][0].sum() > 0
    for k in ("counts", "rendered_cost", "observed_cost", "points_diff_cost"):
        assert np.array_equal(a[k], b[k]), k
