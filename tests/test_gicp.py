"""Batched GICP refinement (the do_icp block, renderer.cu:1688-1817).  fast_gicp is not in the reference tree (parity
unpinned): the CUDA path is held to oracle_gicp, a double-precision restatement of the published
algorithm, within 0.1 mm / 0.01 deg, and to algorithm-level properties (recovers a known transform, does not move a
registered cloud, do_icp == refine-then-score)."""
import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import Oracle
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _rot(axis, ang):
    axis = np.asarray(axis, float) / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def _surface(n=30):
    u, v = np.meshgrid(np.linspace(-.1, .1, n), np.linspace(-.1, .1, n))
    u, v = u.ravel(), v.ravel()
    return np.stack([u, v, 0.8 + 0.03 * np.sin(25 * u) * np.cos(18 * v) + 2 * u * u], 1).astype(np.float32)


def _angle_deg(Ra, Rb):
    """Rotation angle between two rotations from the skew part of Ra^T Rb (sin of the angle): unlike arccos of the
    trace it stays accurate for tiny angles when one matrix is only float32."""
    D = Ra.T @ Rb
    v = 0.5 * np.array([D[2, 1] - D[1, 2], D[0, 2] - D[2, 0], D[1, 0] - D[0, 1]])
    sn, cs = np.linalg.norm(v), (np.trace(D) - 1) / 2
    return np.degrees(np.arctan2(sn, cs))


# ---- CPU: the oracle itself ----------------------------------------------------------------------------------------------
def test_oracle_recovers_known_transform(orc):
    tgt = _surface()
    rng = np.random.default_rng(0)
    R, t = _rot([.3, 1, .2], np.deg2rad(3.0)), np.array([0.004, -0.003, 0.005])
    src = ((tgt[rng.choice(len(tgt), 300, replace=False)] - t) @ R).astype(np.float32)
    T, it, err = orc.gicp(src, tgt)
    assert 1 <= it < 30
    assert _angle_deg(T[:9].reshape(3, 3), R) < 1e-3 and np.abs(T[9:] - t).max() < 1e-6


def test_oracle_identity_cases(orc):
    tgt = _surface(12)
    T, it, _ = orc.gicp(tgt[::3], tgt)          # already registered: one linearisation, no motion
    assert it == 1 and np.allclose(T, [1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0], atol=1e-9)
    T, it, _ = orc.gicp(np.zeros((0, 3), np.float32), tgt)
    assert it == 0 and np.array_equal(T, [1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0])


def test_oracle_label_restriction(orc):
    tgt = np.concatenate([_surface(20), _surface(20) + np.float32([0, 0, 0.02])])
    lab = np.repeat([0, 1], 400).astype(np.int32)
    src = (_surface(20)[::2] + np.float32([0.001, 0, 0.012])).astype(np.float32)   # between both sheets, nearer sheet 1
    T0, _, _ = orc.gicp(src, tgt, lab, 0)
    T1, _, _ = orc.gicp(src, tgt, lab, 1)
    assert T0[11] < -0.008 and T1[11] > 0.004    # pulled onto the sheet of its own label


# ---- GPU ------------------------------------------------------------------------------------------------------------------
def _perturbed(case, n, seed=4, trans=0.006, rot_deg=3.0):
    from perception_b200.renderer import pose_to_mat4x4
    s = case.scene
    rng = np.random.default_rng(seed)
    poses, pmm, lab = [], [], []
    for i in range(n):
        g = i % len(s.gt_poses)
        T = s.gt_poses[g].copy()
        D = np.eye(4)
        D[:3, :3] = _rot(rng.normal(size=3), np.deg2rad(rng.uniform(0, rot_deg)))
        c = T[:3, 3].copy()
        D[:3, 3] = c - D[:3, :3] @ c + rng.uniform(-trans, trans, 3)   # rotate about the object centre, then shift
        poses.append(pose_to_mat4x4(D @ T))
        pmm.append(s.gt_models[g])
        lab.append(g)
    return np.stack(poses), np.array(pmm, np.int32), np.array(lab, np.int32)


@pytest.mark.gpu
@pytest.mark.parametrize("dof6", [False, True])
def test_gicp_matches_oracle(orc, dof6):
    from perception_b200 import synth
    from perception_b200.renderer import Renderer
    case = make_case(2000, 3, 42)
    s = case.scene
    stride = 4
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    poses, pmm, lab = _perturbed(case, 9)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
        r.set_models(s.tris, s.tris_model_count)
        r.set_scene(s.source_depth, s.source_color, s.source_label if dof6 else None)
        r.set_observed(obs["aos"], obs["rgb"], obs["label"] if dof6 else None)
        kw = dict(stride=stride, sensor_resolution=0.01, occlusion_threshold=1.0)
        g = r.gicp(poses, pmm, lab if dof6 else None, **kw)
        g3 = r.gicp(poses, pmm, lab if dof6 else None, icp_max_iterations=3, **kw)
        g1 = r.gicp(poses, pmm, lab if dof6 else None, icp_max_iterations=1, **kw)
        cl = r.render_cuda_multi_unified("CLOUD", poses, pmm, lab if dof6 else None, **kw)
        again = r.gicp(poses, pmm, lab if dof6 else None, **kw)
        assert np.array_equal(g["transforms"], again["transforms"])   # fixed reduction order: deterministic
        # the target exactly as the context holds it: stably sorted by label in 6-DoF (renderer.cu:1681)
        tgt, tlab = obs["aos"], None
        if dof6:
            order = np.argsort(obs["label"], kind="stable")
            tgt, tlab = obs["aos"][order], obs["label"][order].astype(np.int32)
        worst_t = worst_r = 0.0
        converged = moved = same_path = 0
        other_path = []
        for p in range(len(poses)):
            src = cl["result_cloud"][:, cl["result_cloud_pose_map"] == p].T
            assert len(src) > 30
            L = int(lab[p]) if dof6 else 0
            # (a) one Gauss-Newton step from the identical state: comparable for every pose, wild ones included
            T1, it1, _ = orc.gicp(src, tgt, tlab, L, max_iterations=1)
            Tg1 = g1["transforms"][p].astype(np.float64)
            assert int(g1["iterations"][p]) == it1
            worst_t = max(worst_t, np.abs(Tg1[9:] - T1[9:]).max())
            worst_r = max(worst_r, _angle_deg(Tg1[:9].reshape(3, 3), T1[:9].reshape(3, 3)))
            # three steps: a nearest-neighbour near-tie that resolves differently in float32 changes one correspondence
            # of a later step, so wild poses are only held to 0.5 mm / 0.1 deg here (converging ones to the strict
            # tolerance, below)
            T3, it3, _ = orc.gicp(src, tgt, tlab, L, max_iterations=3)
            Tg3 = g3["transforms"][p].astype(np.float64)
            assert int(g3["iterations"][p]) == it3
            d3t, d3r = np.abs(Tg3[9:] - T3[9:]).max(), _angle_deg(Tg3[:9].reshape(3, 3), T3[:9].reshape(3, 3))
            assert d3t < 5e-4 and d3r < 0.1, (p, d3t, d3r)
            # (b) run to convergence with the reference's settings (150 iterations, 2e-3 / 5e-4): comparable where the
            # iteration converges; a Gauss-Newton sequence that keeps switching correspondences is chaotic in the last bits
            T, it, err = orc.gicp(src, tgt, tlab, L)
            if it >= 150:
                continue
            converged += 1
            Tg = g["transforms"][p].astype(np.float64)
            moved += np.abs(T[9:]).max() > 1e-3
            if abs(int(g["iterations"][p]) - it) <= 1:
                # same sequence of linearisations: the end points agree to the tolerance
                same_path += 1
                worst_t, worst_r = max(worst_t, d3t), max(worst_r, d3r)
                worst_t = max(worst_t, np.abs(Tg[9:] - T[9:]).max())
                worst_r = max(worst_r, _angle_deg(Tg[:9].reshape(3, 3), T[:9].reshape(3, 3)))
            else:
                # the float32 sequence took another branch at a correspondence switch: Gauss-Newton with switching
                # correspondences is chaotic in the last bits, and on a badly perturbed pose the two runs can stop
                # centimetres apart.  Step parity is what the three-step comparison above holds for EVERY pose; here such
                # poses are only counted (and bounded below).
                other_path.append((p, int(g["iterations"][p]), it))
            # adjusted pose = T * pose (metres) re-scaled by 100 (renderer.cu:1412-1429)
            M = poses[p].astype(np.float64).reshape(4, 4)
            A = M.copy(); A[:3] /= 100.0
            want = np.eye(4); want[:3, :3] = Tg[:9].reshape(3, 3); want[:3, 3] = Tg[9:]
            want = want @ A; want[:3] *= 100.0
            np.testing.assert_allclose(g["adjusted_poses"][p].reshape(4, 4), want, atol=1e-3, rtol=1e-5)  # cm units
        print(f"GICP vs oracle: max |dt| = {worst_t * 1e3:.4f} mm, max rotation difference = {worst_r:.5f} deg; "
              f"{converged}/{len(poses)} poses converged, {same_path} along the same sequence; others: {other_path}")
        assert worst_t < 1e-4 and worst_r < 0.01      # 0.1 mm / 0.01 deg
        assert converged >= 2 * len(poses) // 3 and moved >= converged // 2 and same_path >= (2 * converged) // 3


@pytest.mark.gpu
def test_do_icp_equals_refine_then_score(orc):
    from perception_b200 import synth
    from perception_b200.renderer import Renderer
    case = make_case(2000, 3, 42)
    s = case.scene
    stride = 4
    obs = observed_6dof(case, stride, orc)
    poses, pmm, lab = _perturbed(case, 12, seed=9)
    tot = np.bincount(obs["label"], minlength=len(s.gt_poses)).astype(np.float32)[lab]
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
        r.set_models(s.tris, s.tris_model_count)
        r.set_scene(s.source_depth, s.source_color, s.source_label)
        r.set_observed(obs["aos"], obs["rgb"], obs["label"])
        kw = dict(stride=stride, cost_type=2, sensor_resolution=0.01, occlusion_threshold=1.0)
        plain = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, **kw)
        icp = r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, do_icp=True, **kw)
        g = r.gicp(poses, pmm, lab, stride=stride, sensor_resolution=0.01, occlusion_threshold=1.0)
        rescored = r.render_cuda_multi_unified("COST", g["adjusted_poses"], pmm, lab, tot, **kw)
        st = r.stats()
    assert np.array_equal(icp["adjusted_poses"], g["adjusted_poses"])
    assert np.array_equal(icp["counts"], rescored["counts"])
    assert np.array_equal(icp["rendered_cost"], rescored["rendered_cost"])
    assert st["icp_runtime"] > 0
    # refinement helps: mean rendered cost of the perturbed ground-truth poses drops
    ok = plain["rendered_cost"] >= 0
    print("rendered cost before/after GICP:", plain["rendered_cost"][ok].mean(), icp["rendered_cost"][ok].mean())
    assert icp["rendered_cost"][ok].mean() < plain["rendered_cost"][ok].mean()


@pytest.mark.gpu
def test_gicp_edge_cases(orc):
    from perception_b200 import synth
    from perception_b200.renderer import Renderer, pose_to_mat4x4
    case = make_case(600, 3, 42)
    s = case.scene
    obs = observed_3dof(case, 4, orc)
    far = s.gt_poses[0].copy(); far[0, 3] += 5.0
    poses = np.stack([pose_to_mat4x4(s.gt_poses[0]), pose_to_mat4x4(far)])
    pmm = np.array([s.gt_models[0]] * 2, np.int32)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
        r.set_models(s.tris, s.tris_model_count)
        r.set_scene(s.source_depth, s.source_color, None)
        r.set_observed(obs["aos"], obs["rgb"], None)
        g = r.gicp(poses, pmm, None, stride=4, sensor_resolution=0.0075)
        assert g["iterations"][1] == 0 and np.array_equal(g["adjusted_poses"][1], poses[1])   # renders nothing: untouched
        assert g["iterations"][0] >= 1
        e = r.gicp(np.zeros((0, 4, 4), np.float32), np.zeros(0, np.int32), None, stride=4)
        assert e["adjusted_poses"].shape == (0, 4, 4)
        r.set_observed(np.zeros((0, 3), np.float32), None, None)                              # no target at all
        g0 = r.gicp(poses[:1], pmm[:1], None, stride=4, sensor_resolution=0.0075)
        assert g0["iterations"][0] == 0 and np.allclose(g0["adjusted_poses"][0], poses[0], atol=1e-4)


@pytest.mark.gpu
def test_gicp_dense_scene_cloud_grids_agree(orc, monkeypatch):
    """BASELINE configs[3] in shape: the target is the FULL-resolution scene cloud (stride 1).  GICP then searches its own
    finer grid, falls back to the cost grid and finally to a linear scan, or searches a pyramid of grids; all are exact, so the
    refined poses must not depend on which one answers.  Forced three ways: fine grid (default), cost grid only, and a fine grid so fine
    that most queries leave it.  (The cost-grid path is the one held to the double-precision oracle by the tests above; the
    oracle's brute-force normals are O(N^2) and out of reach at 200k points.)"""
    from perception_b200 import synth
    from perception_b200.renderer import Renderer
    case = make_case(600, 3, 42)
    s = case.scene
    cam = (case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
    poses, pmm, _ = poses_near_gt(case, 24, trans_sigma=0.01, rot_sigma=0.05)
    far = poses[5].copy(); far[2, 3] -= 25.0   # 25 cm in front of everything: beyond the fine grid's rings
    poses = np.concatenate([poses, far[None]])
    pmm = np.concatenate([pmm, pmm[5:6]])
    outs = []
    # last entry: target normals from the N x N k-NN kernel instead of the ring search through the grid (same 10 neighbours
    # in the same order => bit-identical normals => identical poses)
    # 4th field: PERCH_GICP_PYRAMID (1 = default: pyramid of grids for sparse targets only; 2 = forced on; 0 = off)
    for density, target, brute, pyramid in (("8", "60", "0", "1"), ("1000000", "60", "0", "1"), ("8", "3", "0", "1"), ("8", "60", "1", "1"),
                                            ("8", "200", "0", "1"), ("8", "60", "0", "2"), ("1000000", "60", "0", "0")):
        monkeypatch.setenv("PERCH_GICP_GRID_DENSITY", density)
        monkeypatch.setenv("PERCH_GICP_GRID_TARGET_X10", target)
        monkeypatch.setenv("PERCH_GICP_NORMALS_BRUTE", brute)
        monkeypatch.setenv("PERCH_GICP_PYRAMID", pyramid)
        with Renderer(0) as r:
            r.set_camera(case.width, case.height, case.proj, *cam)
            r.set_models(s.tris, s.tris_model_count)
            r.set_scene(s.source_depth, s.source_color, None)
            obs = r.depth2cloud_global(s.source_depth, s.source_color, case.width, case.height, 1, *cam, 1)
            assert obs["rendered_cloud_point_num"] > 150000
            r.set_observed(obs["result_cloud_eigen"], obs["result_cloud_color"], None)
            outs.append(r.gicp(poses, pmm, None, stride=4, sensor_resolution=0.0075, icp_max_iterations=8))
    assert outs[0]["iterations"].max() >= 2
    for o in outs[1:]:
        assert np.array_equal(o["iterations"], outs[0]["iterations"])
        assert np.array_equal(o["transforms"], outs[0]["transforms"])


@pytest.mark.gpu
@pytest.mark.parametrize("dof6", [False, True])
def test_gicp_converges_before_the_iteration_cap(orc, dof6):
    """With the Levenberg-Marquardt step (fast_gicp's default optimiser) strongly perturbed poses settle instead of
    switching correspondences until the 150-iteration cap (round 1: plain Gauss-Newton, about a third never stopped):
    >= 95 % of 60 perturbed poses must stop before the cap at the reference's settings (renderer.cu:1696-1699), and the
    refinement must actually register them (lower final error than after the first linearisation)."""
    from perception_b200 import synth
    from perception_b200.renderer import Renderer
    case = make_case(2000, 3, 42)
    s = case.scene
    stride = 4
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    poses, pmm, lab = _perturbed(case, 60, seed=9)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
        r.set_models(s.tris, s.tris_model_count)
        r.set_scene(s.source_depth, s.source_color, s.source_label if dof6 else None)
        r.set_observed(obs["aos"], obs["rgb"], obs["label"] if dof6 else None)
        kw = dict(stride=stride, sensor_resolution=0.01, occlusion_threshold=1.0)
        g1 = r.gicp(poses, pmm, lab if dof6 else None, icp_max_iterations=1, **kw)
        g = r.gicp(poses, pmm, lab if dof6 else None, **kw)   # reference settings: 150 iterations, 2e-3 / 5e-4
    it = g["iterations"]
    stopped = (it < 150).mean()
    print(f"GICP iterations: median {np.median(it)}, max {it.max()}, stopped before the cap: {100 * stopped:.1f} %")
    assert stopped >= 0.95
    assert (g["error"] <= g1["error"] * 1.0001 + 1e-12).mean() >= 0.9
    assert np.median(it) < 40
