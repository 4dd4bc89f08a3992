"""Caller glue: ObjectState lists -> pose matrices / observed totals (GetStateImagesUnifiedGPU,
search_env.cpp:1505-1622), IsValidPose (309-528), 3-DoF GenerateSuccessorStates (7268-7472).

Eigen / PCL / FLANN are not in the reference tree (parity unpinned at that boundary, oracle/perch_states_oracle.c): the
CUDA path is held bit for bit to the CPU restatement -- integers, validity flags and generated lists exactly; the float
pose matrices exactly up to the device's double sin/cos (<= 1 float ulp tolerated, counted) -- and the restatement
itself to a numpy double evaluation of the same formulas."""
import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import StatesOracle
from perception_b200 import states as S
from states_cases import CAM_TO_BODY, case_3dof, case_6dof, states_3dof, states_6dof


def _rot_xyz(roll, pitch, yaw):
    cr, sr, cp, sp, cy, sy = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _numpy_pose(env, meta, st):
    """(cam_to_world * cam_to_body)^-1 * GetTransform() * preprocessing_transform in numpy doubles."""
    T = np.eye(4)
    if st["qx"] == 0 and st["qy"] == 0 and st["qz"] == 0 and st["qw"] == 0:
        T[:3, :3] = _rot_xyz(st["roll"], st["pitch"], st["yaw"])
    else:
        w, x, y, z = np.array([st["qw"], st["qx"], st["qy"], st["qz"]]) / np.linalg.norm([st["qw"], st["qx"], st["qy"], st["qz"]])
        T[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
    T[:3, 3] = [st["x"], st["y"], st["z"]]
    P = np.linalg.inv(np.asarray(env.cam_to_world) @ CAM_TO_BODY) @ T @ np.asarray(meta.preprocessing_transform, np.float64)
    P[:3] *= 100
    return P


def _ulps(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    scale = np.spacing(np.maximum(np.abs(a), np.abs(b)).astype(np.float32))
    return np.abs(a.astype(np.float64) - b.astype(np.float64)) / scale


# ---- CPU: the restatement against numpy ---------------------------------------------------------------------------------
def test_oracle_camera_matrix_and_poses():
    c = case_3dof()
    orc = StatesOracle(c["env"], c["metas"], c["projected"])
    cam = orc.camera_matrix()
    assert np.allclose(cam, np.linalg.inv(np.asarray(c["env"].cam_to_world) @ CAM_TO_BODY), atol=1e-14)
    st = states_3dof(c)
    poses, pmm, lab, tot = orc.assemble(st, observed_point_num=len(c["projected"]))
    assert np.array_equal(pmm, st["model_id"]) and np.all(lab == -1)
    for i in range(len(st)):
        ref = _numpy_pose(c["env"], c["metas"][st["model_id"][i]], st[i])
        assert np.allclose(poses[i], ref, rtol=0, atol=1e-4), i       # cm units: 1e-4 cm
        assert _ulps(poses[i][np.abs(ref) > 1e-3], ref[np.abs(ref) > 1e-3].astype(np.float32)).max() <= 1
    # the scene's own placement convention gives back the ground-truth camera poses
    s = c["case"].scene
    for i, T in enumerate(s.gt_poses):
        assert np.allclose(poses[i][:3] / 100, T[:3], atol=1e-6)
    # cylinder totals: brute force in numpy with the same float expression
    xy = c["projected"]
    for i in range(0, len(st), 7):
        m = c["metas"][st["model_id"][i]]
        r2 = np.float32((m.inflation_factor * m.circumscribed_radius) ** 2)
        q = np.float32([st["x"][i], st["y"][i], st["z"][i]])
        d = (q[0] - xy[:, 0]) ** 2
        d = d + (q[1] - xy[:, 1]) ** 2
        d = d + (q[2] - xy[:, 2]) ** 2
        assert tot[i] == np.count_nonzero(d < r2)


def test_oracle_6dof_labels_and_totals():
    c = case_6dof()
    orc = StatesOracle(c["env"], c["metas"], c["xyz"], c["ranges"])
    st = states_6dof(c)
    seg = np.bincount(c["obs"]["label"], minlength=c["n_labels"]).astype(np.float32)
    poses, pmm, lab, tot = orc.assemble(st, seg_counts=seg)
    assert np.array_equal(lab, st["segmentation_label_id"] - 1)
    assert np.array_equal(tot, seg[lab])
    for i in range(0, len(st), 5):
        ref = _numpy_pose(c["env"], c["metas"][pmm[i]], st[i])
        assert np.allclose(poses[i], ref, rtol=0, atol=2e-4)
    valid = orc.validate(st)
    assert 0 < valid.sum() < len(st)   # near-ground-truth candidates pass, the far-off ones do not
    # validity is the segmented cloud's neighbour count
    for i in range(0, len(st), 9):
        m = c["metas"][pmm[i]]
        b, e = c["ranges"][1 + lab[i]]
        rad = max(m.inflation_factor * m.circumscribed_radius_3d, np.hypot(c["env"].res / 2, c["env"].res / 2))
        q = np.float32([st["x"][i], st["y"][i], st["z"][i]])
        d = (q[0] - c["xyz"][b:e, 0]) ** 2
        d = d + (q[1] - c["xyz"][b:e, 1]) ** 2
        d = d + (q[2] - c["xyz"][b:e, 2]) ** 2
        assert valid[i] == (np.count_nonzero(d < np.float32(rad * rad)) >= c["env"].min_neighbor_points_for_valid_pose)


def test_oracle_generate_rules():
    c = case_3dof(symmetric=(False, True, False), symmetry_mode=(0, 0, 1))
    orc = StatesOracle(c["env"], c["metas"], c["projected"])
    gen = orc.generate()
    assert len(gen) > 50
    # reference order: model, then x, then y, then yaw
    key = np.stack([gen["model_id"], gen["x"], gen["y"], gen["yaw"]], 1)
    assert all(tuple(key[i]) < tuple(key[i + 1]) for i in range(len(key) - 1))
    assert orc.validate(gen).all()
    # the symmetric model keeps one yaw per cell, the semi-symmetric one stops above pi + theta_res
    sym = gen[gen["model_id"] == 1]
    assert len(np.unique(np.stack([sym["x"], sym["y"]], 1), axis=0)) == len(sym)
    semi = gen[gen["model_id"] == 2]
    assert semi["yaw"].max() <= np.pi + c["env"].theta_res + 1e-12 and semi["yaw"].max() > np.pi - 1e-9
    # grid values are the accumulated doubles of `for (x = x_min; x <= x_max; x += res)`
    xs, x = [], c["env"].x_min
    while x <= c["env"].x_max:
        xs.append(x)
        x += c["env"].res
    assert set(np.unique(gen["x"])) <= set(xs)
    # placed objects: their model generates nothing; others keep clear of the inscribed circles
    gen0 = gen[gen["model_id"] == 0]
    placed = gen0[len(gen0) // 2:][:1]
    gen2 = orc.generate(placed)
    assert not np.any(gen2["model_id"] == 0) and 0 < len(gen2) < len(gen[gen["model_id"] != 0])
    r0 = c["metas"][0].inscribed_radius
    for m in (1, 2):
        g = gen2[gen2["model_id"] == m]
        d2 = (g["x"] - placed["x"][0]) ** 2 + (g["y"] - placed["y"][0]) ** 2
        assert np.all(d2 >= (r0 + c["metas"][m].inscribed_radius) ** 2)


def test_oracle_constraint_cloud():
    c = case_3dof(with_constraint=True, min_neighbor=0)
    free = StatesOracle(c["env"], c["metas"], c["projected"]).generate()
    tied = StatesOracle(c["env"], c["metas"], c["projected"], constraint_xyz=c["constraint"]).generate()
    assert 0 < len(tied) < len(free)
    # every surviving pose keeps the constraint points' centroid within its circumscribed radius
    cen = c["constraint"][:, :2].mean(0)
    for g in tied:
        m = c["metas"][g["model_id"]]
        assert np.hypot(g["x"] - cen[0], g["y"] - cen[1]) <= m.inflation_factor * m.circumscribed_radius + 0.03


def test_state_layout_matches_header():
    import ctypes as C
    from perception_b200._capi import perch_env, perch_model_meta, perch_state
    assert C.sizeof(perch_state) == S.STATE_DTYPE.itemsize == 88
    assert C.sizeof(perch_model_meta) == 64 + 5 * 8 + 4 * 4
    assert C.sizeof(perch_env) == 16 * 8 + 8 * 8 + 6 * 4
    for name in S.STATE_DTYPE.names:
        assert S.STATE_DTYPE.fields[name][1] == getattr(perch_state, name).offset


# ---- GPU: parity of the CUDA glue with the restatement --------------------------------------------------------------------
gpu = pytest.mark.gpu


def _renderer_3dof(c, constraint=False):
    from perception_b200.renderer import Renderer
    case, s = c["case"], c["case"].scene
    r = Renderer(0)
    r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, 100.0)
    r.set_models(s.tris, s.tris_model_count)
    r.set_scene(s.source_depth, s.source_color)
    r.set_observed(c["obs"]["aos"], c["obs"]["rgb"])
    r.set_env(c["env"], c["metas"])
    r.set_search_cloud(c["projected"], constraint_xyz=c["constraint"] if constraint else None)
    return r


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_assemble_parity_3dof():
    c = case_3dof()
    st = states_3dof(c, n=500)
    orc = StatesOracle(c["env"], c["metas"], c["projected"])
    o_pose, o_pmm, o_lab, o_tot = orc.assemble(st, observed_point_num=len(c["obs"]["aos"]))
    with _renderer_3dof(c) as r:
        g_pose, g_pmm, g_lab, g_tot = r.assemble_states(st)
        assert np.array_equal(g_pmm, o_pmm) and np.array_equal(g_lab, o_lab)
        assert np.array_equal(g_tot, o_tot)                       # integer counts
        exact = g_pose == o_pose
        assert exact.mean() > 0.999, exact.mean()
        big = np.abs(o_pose) > 1e-3
        assert _ulps(g_pose[big], o_pose[big]).max() <= 1
        assert np.abs(g_pose - o_pose).max() < 1e-5
        # cylinder off: every pose gets observed_point_num (search_env.cpp:1615-1619)
        env2 = S.Env(**{**c["env"].__dict__, "use_cylinder_observed": False})
        r.set_env(env2, c["metas"])
        assert np.all(r.assemble_states(st)[3] == len(c["obs"]["aos"]))


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_assemble_and_validate_parity_6dof():
    from perception_b200.renderer import Renderer
    c = case_6dof()
    case, s = c["case"], c["case"].scene
    st = states_6dof(c, n=400)
    orc = StatesOracle(c["env"], c["metas"], c["xyz"], c["ranges"])
    seg = np.bincount(c["obs"]["label"], minlength=c["n_labels"]).astype(np.float32)
    o_pose, o_pmm, o_lab, o_tot = orc.assemble(st, seg_counts=seg)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, 100.0)
        r.set_models(s.tris, s.tris_model_count)
        r.set_scene(s.source_depth, s.source_color, s.source_label)
        r.set_observed(c["obs"]["aos"], c["obs"]["rgb"], c["obs"]["label"])
        r.set_env(c["env"], c["metas"])
        r.set_search_cloud(c["xyz"], c["ranges"])
        g_pose, g_pmm, g_lab, g_tot = r.assemble_states(st)
        assert np.array_equal(g_pmm, o_pmm) and np.array_equal(g_lab, o_lab) and np.array_equal(g_tot, o_tot)
        assert np.array_equal(g_pose, o_pose)                     # quaternion path: no sin/cos, bit for bit
        assert np.array_equal(r.validate_states(st), orc.validate(st))
        assert np.array_equal(r.validate_states(st, after_refinement=True), orc.validate(st, after_refinement=True))
        # one call == assemble + render_cuda_multi_unified
        kw = dict(stride=4, cost_type=2, sensor_resolution=0.01)
        a = r.score_states("COST", st, **kw)
        b = r.render_cuda_multi_unified("COST", g_pose, g_pmm, g_lab, g_tot, **kw)
        for k in ("rendered_cost", "observed_cost", "points_diff_cost", "counts"):
            assert np.array_equal(a[k], b[k], equal_nan=True), k
        # a label without observed points is an error, not an out-of-bounds read
        bad = st[:4].copy()
        bad["segmentation_label_id"] = 40
        with pytest.raises(Exception, match="segmentation_label_id"):
            r.assemble_states(bad)
        bad = st[:4].copy()
        bad["model_id"] = 17
        with pytest.raises(Exception, match="model_id"):
            r.score_states("COST", bad, **kw)


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
@pytest.mark.parametrize("constraint", [False, True])
def test_validate_and_generate_parity_3dof(constraint):
    c = case_3dof(with_constraint=constraint, symmetric=(False, True, False), symmetry_mode=(0, 0, 1), res=0.03,
                  theta_res=np.pi / 8)
    st = states_3dof(c, n=600)
    orc = StatesOracle(c["env"], c["metas"], c["projected"], constraint_xyz=c["constraint"] if constraint else None)
    with _renderer_3dof(c, constraint) as r:
        gen_o = orc.generate()
        gen_g = r.generate_states()
        assert len(gen_g) == len(gen_o) > 0
        assert gen_g.tobytes() == gen_o.tobytes()                 # same states, same order, every bit
        placed = gen_o[[0, len(gen_o) // 2]]
        ov, gv = orc.validate(st, placed), r.validate_states(st, placed)
        assert np.array_equal(gv, ov) and 0 < ov.sum() < len(st)
        gen2_o, gen2_g = orc.generate(placed), r.generate_states(placed)
        assert gen2_g.tobytes() == gen2_o.tobytes() and len(gen2_o) < len(gen_o)


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_generate_then_score_without_host_round_trip():
    c = case_3dof(res=0.04, theta_res=np.pi / 8)
    with _renderer_3dof(c) as r:
        gen = r.generate_states()
        kw = dict(stride=4, cost_type=0, sensor_resolution=0.0075)
        dev = r.score_states("COST", None, **kw)                 # the list left on the device
        host = r.score_states("COST", gen, **kw)                 # the same list sent from the host
        poses, pmm, lab, tot = r.assemble_states(gen)
        ref = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, **kw)
        for k in ("rendered_cost", "observed_cost", "points_diff_cost", "counts"):
            assert np.array_equal(dev[k], host[k], equal_nan=True) and np.array_equal(dev[k], ref[k], equal_nan=True), k
        assert len(dev["rendered_cost"]) == len(gen) == r.n_generated
        # the best candidate of each model sits at its ground-truth placement (grid pitch + yaw step away at most)
        from perception_b200.renderer import topk
        hits = topk(dev["rendered_cost"], dev["observed_cost"], pmm, len(c["metas"]), 1)
        s = c["case"].scene
        T_table_cam = np.linalg.inv(s.T_cam_table)
        for m in set(s.gt_models):
            h = hits[m][0]
            assert h["pose_id"] >= 0
            best = gen[h["pose_id"]]
            d = min(np.hypot(best["x"] - (T_table_cam @ T)[0, 3], best["y"] - (T_table_cam @ T)[1, 3])
                    for T, gm in zip(s.gt_poses, s.gt_models) if gm == m)
            assert d <= 0.04


@gpu
@pytest.mark.skipif(not has_gpu(), reason="needs a B200")
def test_states_call_order_errors():
    from perception_b200.renderer import Renderer
    c = case_3dof()
    with Renderer(0) as r:
        with pytest.raises(Exception, match="perch_set_env"):
            r.assemble_states(states_3dof(c, n=4))
        r.set_env(c["env"], c["metas"])
        with pytest.raises(Exception, match="perch_set_search_cloud"):
            r.generate_states()
        with pytest.raises(Exception, match="cylinder"):
            r.assemble_states(states_3dof(c, n=4))
