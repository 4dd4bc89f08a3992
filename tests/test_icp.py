"""cuda_icp package API (point-to-plane ICP against a depth image, projective association): oracle pinned against the
reference's own sources (scene preparation bit-exact vs its host code; ICP loop vs its GPU loop, to float-accumulation
accuracy because the reference's reduction order is unspecified), product held to the oracle."""
from pathlib import Path

import numpy as np
import pytest

import icp_cases
from conftest import has_gpu
from oracle_lib import ICP_REF_SO, Oracle, ReferenceIcp

GOLDEN = Path(__file__).parent / "golden" / "icp.npz"
TOL_T, TOL_DEG = 1e-4, 0.01   # 0.1 mm / 0.01 deg (class 4)


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _angle_deg(Ra, Rb):
    D = np.asarray(Ra, np.float64).T @ np.asarray(Rb, np.float64)
    v = 0.5 * np.array([D[2, 1] - D[1, 2], D[0, 2] - D[2, 0], D[1, 0] - D[0, 1]])
    return np.degrees(np.arctan2(np.linalg.norm(v), (np.trace(D) - 1) / 2))


def _close(Ta, Tb):
    return np.abs(Ta[:3, 3] - Tb[:3, 3]).max() < TOL_T and _angle_deg(Ta[:3, :3], Tb[:3, :3]) < TOL_DEG


# ---- CPU -------------------------------------------------------------------------------------------------------------------
def test_oracle_scene_matches_reference_host_code(orc):
    case, K = icp_cases.scene()
    pcd, nrm = orc.icp_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K)
    assert (np.abs(nrm).sum(1) > 0).sum() > 1000
    np.testing.assert_allclose(np.linalg.norm(nrm[np.abs(nrm).sum(1) > 0], axis=1), 1.0, atol=1e-5)
    if GOLDEN.exists():   # digest of the buffers the reference's host code produced on the golden run
        import hashlib
        digest = np.frombuffer(hashlib.sha256(pcd.tobytes() + nrm.tobytes()).digest(), np.uint8)
        assert np.array_equal(digest, np.load(GOLDEN)["scene_sha256"])
    if ICP_REF_SO.exists():
        try:
            ref = ReferenceIcp()
        except OSError as e:
            pytest.skip(f"cannot load the reference cuda_icp library: {e}")
        r_pcd, r_nrm = ref.scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K)   # host code: no GPU needed
        assert np.array_equal(pcd, r_pcd) and np.array_equal(nrm, r_nrm)


def test_oracle_icp_undoes_a_known_motion(orc):
    case, K = icp_cases.scene()
    pcd, nrm = orc.icp_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K)
    cl, moves = icp_cases.clouds(pcd, case.scene.source_label)
    exact = 0
    for c, (R, t) in zip(cl, moves):
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K)
        before = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_iteration=0)
        assert o["fitness"] > 0.9 and o["inlier_rmse"] < before["inlier_rmse"]   # point-to-plane error went down
        assert o["iterations"] >= 1
        M = np.eye(4); M[:3, :3] = R; M[:3, 3] = t
        res = o["transformation"].astype(np.float64) @ M                            # ICP undoes the motion
        exact += np.abs(res - np.eye(4)).max() < 1e-5
    assert exact >= len(cl) // 2
    # empty cloud / no valid correspondence: identity, zeros (icp.cu:186)
    z = orc.icp_point2plane(np.zeros((0, 3), np.float32), pcd, nrm, icp_cases.W, icp_cases.H, K)
    assert np.array_equal(z["transformation"], np.eye(4)) and z["fitness"] == 0
    far = orc.icp_point2plane(cl[0] + np.float32([0, 0, 5]), pcd, nrm, icp_cases.W, icp_cases.H, K)
    assert np.array_equal(far["transformation"], np.eye(4)) and far["fitness"] == 0


def test_oracle_icp_matches_reference_golden(orc):
    if not GOLDEN.exists():
        pytest.skip("tests/golden/icp.npz not generated yet")
    g = np.load(GOLDEN)
    case, K = icp_cases.scene()
    pcd, nrm = orc.icp_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K)
    cl, _ = icp_cases.clouds(pcd, case.scene.source_label)
    converged = 0
    for i, c in enumerate(cl):
        # After a fixed number of updates the two runs need not agree to 0.1 mm: the association is discrete (a point
        # whose projection sits within rounding of a pixel border picks another scene pixel), so a last-bit difference in
        # the float sums -- the reference's reduction order is unspecified -- shifts the path; only the quality is held.
        o3 = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_iteration=3)
        assert abs(o3["fitness"] - float(g[f"fit3_{i}"])) < 0.02
        assert abs(o3["inlier_rmse"] - float(g[f"rmse3_{i}"])) < 0.2 * float(g[f"rmse3_{i}"]) + 1e-6
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K)
        if o["iterations"] >= 30:
            continue   # runs to the cap while correspondences keep flipping: chaotic in the last bits, not comparable
        converged += 1
        assert _close(o["transformation"], g[f"T{i}"]), i
        assert abs(o["fitness"] - float(g[f"fit{i}"])) < 5e-3 and abs(o["inlier_rmse"] - float(g[f"rmse{i}"])) < 1e-5
    assert converged >= 2


def test_oracle_icp_nn_matches_reference_golden(orc):
    """Scene_nn association (pcd_scene.h:60-136): golden = the reference's own kd-tree walk inside
    ICP_Point2Plane_cuda<Scene_nn> (tests/golden/make_golden_icp.py); the oracle finds the same correspondences by brute
    force."""
    path = GOLDEN.parent / "icp_nn.npz"
    if not path.exists():
        pytest.skip("tests/golden/icp_nn.npz not generated yet")
    g = np.load(path)
    case, K = icp_cases.scene()
    pcd, nrm = orc.icp_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K)
    cl, _ = icp_cases.clouds(pcd, case.scene.source_label)
    for i, c in enumerate(cl[:2]):   # the brute-force association is O(points x pixels) per iteration
        o0 = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_dist_diff=0.01, max_iteration=0, nn=True)
        assert abs(o0["fitness"] - float(g[f"fit0_{i}"])) < 1e-6 and abs(o0["inlier_rmse"] - float(g[f"rmse0_{i}"])) < 1e-6
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_dist_diff=0.01, nn=True)
        if o["iterations"] < 30:
            assert _close(o["transformation"], g[f"T{i}"]), i
            assert abs(o["fitness"] - float(g[f"fit{i}"])) < 5e-3 and abs(o["inlier_rmse"] - float(g[f"rmse{i}"])) < 1e-5


# ---- GPU -------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_icp_scene_bit_exact(orc):
    from perception_b200.renderer import Renderer
    case, K = icp_cases.scene()
    with Renderer(0) as r:
        pcd, nrm = r.icp_set_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K, want_buffers=True)
    o_pcd, o_nrm = orc.icp_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K)
    assert np.array_equal(pcd, o_pcd)
    assert np.array_equal(nrm, o_nrm), f"{np.count_nonzero(np.any(nrm != o_nrm, axis=1))} normals differ"


@pytest.mark.gpu
def test_icp_point2plane_parity(orc):
    from perception_b200.renderer import Renderer
    case, K = icp_cases.scene()
    with Renderer(0) as r:
        pcd, nrm = r.icp_set_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K, want_buffers=True)
        cl, _ = icp_cases.clouds(pcd, case.scene.source_label, n_clouds=6)
        batch = r.icp_point2plane(cl + [np.zeros((0, 3), np.float32), cl[0] + np.float32([0, 0, 5])])
        single = [r.icp_point2plane([c]) for c in cl]
        few = r.icp_point2plane(cl, max_iteration=2)
    ref = ReferenceIcp() if ICP_REF_SO.exists() else None
    for i, c in enumerate(cl):
        T = batch["transformation"][i]
        assert np.array_equal(T, single[i]["transformation"][0])           # clouds of a batch are independent
        # fixed number of updates: quality only (discrete association, see test_oracle_icp_matches_reference_golden)
        o2 = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_iteration=2)
        assert few["iterations"][i] == o2["iterations"] == 2
        assert abs(few["fitness"][i] - o2["fitness"]) < 0.02
        assert abs(few["inlier_rmse"][i] - o2["inlier_rmse"]) < 0.2 * o2["inlier_rmse"] + 1e-6
        if ref is not None:   # the reference's own GPU loop, live
            q2 = ref.point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_iteration=2)
            assert abs(few["fitness"][i] - q2["fitness"]) < 0.02
            assert abs(few["inlier_rmse"][i] - q2["inlier_rmse"]) < 0.2 * q2["inlier_rmse"] + 1e-6
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K)
        if o["iterations"] >= 30:
            continue   # never settles (correspondences keep flipping): chaotic in the last bits
        assert _close(T, o["transformation"]), (i, T, o["transformation"])
        assert abs(batch["fitness"][i] - o["fitness"]) < 5e-3 and abs(batch["inlier_rmse"][i] - o["inlier_rmse"]) < 1e-5
        np.testing.assert_allclose(batch["clouds"][i], o["cloud"], atol=2e-4)
        if ref is not None:
            q = ref.point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K)
            assert _close(T, q["transformation"]), (i, T, q["transformation"])
            assert abs(batch["fitness"][i] - q["fitness"]) < 5e-3 and abs(batch["inlier_rmse"][i] - q["inlier_rmse"]) < 1e-5
    n = len(cl)
    assert np.array_equal(batch["transformation"][n], np.eye(4)) and batch["fitness"][n] == 0        # empty cloud
    assert np.array_equal(batch["transformation"][n + 1], np.eye(4)) and batch["fitness"][n + 1] == 0  # nothing valid


@pytest.mark.gpu
def test_icp_nn_parity(orc):
    """ICP_Point2Plane_cuda<Scene_nn> (kd-tree association, pcd_scene.h:47-136): the association itself (fitness and
    rmse before any update) must match the reference's kd-tree walk and the oracle's brute force; settled runs agree to
    0.1 mm / 0.01 deg like the projective variant."""
    from perception_b200.renderer import Renderer
    case, K = icp_cases.scene()
    with Renderer(0) as r:
        pcd, nrm = r.icp_set_scene(case.scene.source_depth, icp_cases.W, icp_cases.H, K, want_buffers=True)
        cl, _ = icp_cases.clouds(pcd, case.scene.source_label, n_clouds=4)
        extra = [np.zeros((0, 3), np.float32), cl[0] + np.float32([0, 0, 5])]
        first = r.icp_point2plane(cl + extra, max_iteration=0, nn_max_dist=0.01)
        full = r.icp_point2plane(cl + extra, nn_max_dist=0.01)
        wide = r.icp_point2plane(cl, max_iteration=0, nn_max_dist=0.03)   # another gate: the grid is rebuilt
        proj = r.icp_point2plane(cl, max_iteration=0)
    ref = ReferenceIcp() if ICP_REF_SO.exists() else None
    g = np.load(GOLDEN.parent / "icp_nn.npz") if (GOLDEN.parent / "icp_nn.npz").exists() else None
    settled = 0
    for i, c in enumerate(cl):
        o0 = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_dist_diff=0.01, max_iteration=0, nn=True)
        assert first["fitness"][i] == np.float32(o0["fitness"]) > 0.5, i          # same correspondences, counted
        assert abs(first["inlier_rmse"][i] - o0["inlier_rmse"]) < 1e-6
        assert wide["fitness"][i] >= first["fitness"][i]
        if g is not None:
            assert abs(first["fitness"][i] - float(g[f"fit0_{i}"])) < 1e-6 and abs(first["inlier_rmse"][i] - float(g[f"rmse0_{i}"])) < 1e-6
        if ref is not None:
            q0 = ref.point2plane_nn(c, case.scene.source_depth, icp_cases.W, icp_cases.H, K, max_iteration=0)
            assert abs(first["fitness"][i] - q0["fitness"]) < 1e-6 and abs(first["inlier_rmse"][i] - q0["inlier_rmse"]) < 1e-6
        o = orc.icp_point2plane(c, pcd, nrm, icp_cases.W, icp_cases.H, K, max_dist_diff=0.01, nn=True)
        if o["iterations"] >= 30:
            continue
        settled += 1
        assert _close(full["transformation"][i], o["transformation"]), i
        assert abs(full["fitness"][i] - o["fitness"]) < 5e-3 and abs(full["inlier_rmse"][i] - o["inlier_rmse"]) < 1e-5
        if ref is not None:
            q = ref.point2plane_nn(c, case.scene.source_depth, icp_cases.W, icp_cases.H, K)
            assert _close(full["transformation"][i], q["transformation"]), i
    assert settled >= 2
    n = len(cl)
    assert np.array_equal(full["transformation"][n], np.eye(4)) and full["fitness"][n] == 0
    assert np.array_equal(full["transformation"][n + 1], np.eye(4)) and full["fitness"][n + 1] == 0
    assert proj["fitness"].shape == (n,)


# ---- the C++ backend for the reference's unchanged <cuda_icp/icp.h> (perception_b200/csrc/compat/cuda_icp_compat.cpp) -------
ICP_COMPAT_SO = Path(__file__).resolve().parent.parent / "oracle" / "_ref" / "libcuda_icp_compat.so"


def _compat():
    import ctypes as C
    if not ICP_COMPAT_SO.exists():
        pytest.skip("oracle/_ref/libcuda_icp_compat.so not built (needs the reference headers)")
    from perception_b200 import _capi
    _capi.load_library()  # libnccl preload order, see _capi._preload_nccl
    lib = C.CDLL(str(ICP_COMPAT_SO))
    lib.compat_icp_last_error.restype = C.c_char_p
    return lib


def _vp(a):
    import ctypes as C
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.gpu
def test_cuda_icp_cpp_signatures_projective_and_nn(orc):
    """cuda_icp::ICP_Point2Plane_cuda<Scene_projective | Scene_nn> and Scene_*::init_*_cuda called through the
    reference's own declarations (device_vector_holder<Vec3f>, cv::Mat, Mat3x3f) == the C-ABI on the same inputs."""
    import ctypes as C
    from perception_b200.renderer import Renderer
    lib = _compat()
    case, K = icp_cases.scene()
    W, H = icp_cases.W, icp_cases.H
    depth = np.ascontiguousarray(case.scene.source_depth, np.int32).reshape(-1)
    K9 = np.ascontiguousarray(K, np.float32).reshape(9)
    with Renderer(0) as r:
        pcd, nrm = r.icp_set_scene(depth, W, H, K, want_buffers=True)
        cl, _ = icp_cases.clouds(pcd, case.scene.source_label, n_clouds=3)
        want_p = r.icp_point2plane(cl)
        want_n = r.icp_point2plane(cl, nn_max_dist=0.01)
    n_valid = int((depth > 0).sum())
    for i, c in enumerate(cl):
        pts = np.ascontiguousarray(c, np.float32).copy()
        T = np.zeros(16, np.float32); fit = C.c_float(0); rmse = C.c_float(0)
        s_pcd, s_nrm = np.zeros((W * H, 3), np.float32), np.zeros((W * H, 3), np.float32)
        rc = lib.compat_icp_projective(_vp(pts), len(pts), _vp(depth), W, H, _vp(K9), C.c_float(0.1), C.c_float(1e-5), C.c_float(1e-5),
                                       30, _vp(T), C.byref(fit), C.byref(rmse), _vp(s_pcd), _vp(s_nrm))
        assert rc == 0, lib.compat_icp_last_error()
        assert np.array_equal(s_pcd, pcd) and np.array_equal(s_nrm, nrm)            # scene prepared on the device
        assert np.array_equal(T.reshape(4, 4), want_p["transformation"][i])
        assert fit.value == want_p["fitness"][i] and rmse.value == want_p["inlier_rmse"][i]
        assert np.array_equal(pts, want_p["clouds"][i])                              # model_pcd transformed in place
        pts = np.ascontiguousarray(c, np.float32).copy()
        n_scene = C.c_int(0)
        rc = lib.compat_icp_nn(_vp(pts), len(pts), _vp(depth), W, H, _vp(K9), C.c_float(1e-5), C.c_float(1e-5), 30, _vp(T),
                               C.byref(fit), C.byref(rmse), C.byref(n_scene))
        assert rc == 0, lib.compat_icp_last_error()
        assert n_scene.value == n_valid
        assert np.array_equal(T.reshape(4, 4), want_n["transformation"][i])
        assert fit.value == want_n["fitness"][i] and rmse.value == want_n["inlier_rmse"][i]


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [32, 16])
def test_cuda_icp_depth2cloud_cuda(bits):
    """cuda_icp::depth2cloud_cuda<int32_t | uint16_t> (icp.h:83-90, icp.cu:225-297) through the reference's declaration:
    == the reference's own kernels at stride 1 (live, oracle/_ref/libicp_ref.so), == the float32 formula at any stride."""
    import ctypes as C
    lib = _compat()
    case, K = icp_cases.scene()
    W, H = icp_cases.W, icp_cases.H
    d32 = np.ascontiguousarray(case.scene.source_depth, np.int32).reshape(H, W)
    depth = d32 if bits == 32 else np.clip(d32, 0, 65535).astype(np.uint16)
    K9 = np.ascontiguousarray(K, np.float32).reshape(9)
    ref = None
    if ICP_REF_SO.exists():
        ref = C.CDLL(str(ICP_REF_SO))
    for stride, tl_x, tl_y in [(1, 0, 0), (1, 17, 5), (4, 0, 0), (8, 3, 2)]:
        n = C.c_int(0)
        cap = W * H
        out = np.zeros((cap, 3), np.float32)
        rc = lib.compat_icp_depth2cloud(_vp(depth), bits, W, H, _vp(K9), stride, tl_x, tl_y, _vp(out), cap, C.byref(n))
        assert rc == 0, lib.compat_icp_last_error()
        sub = depth[::stride, ::stride]
        ys, xs = np.nonzero(sub > 0)                       # row-major order of the lattice
        assert n.value == len(xs) > 0
        z = sub[ys, xs].astype(np.float32) / np.float32(100.0)
        x = ((xs + tl_x).astype(np.float32) - K9[2]) / K9[0] * z
        y = ((ys + tl_y).astype(np.float32) - K9[5]) / K9[4] * z
        assert np.array_equal(out[:n.value], np.stack([x, y, z], 1))
        if ref is not None and stride == 1:
            q = np.zeros((cap, 3), np.float32)
            qn = C.c_int(0)
            assert ref.ref_icp_depth2cloud(_vp(depth), bits, W, H, _vp(K9), stride, tl_x, tl_y, _vp(q), cap, C.byref(qn)) == 0
            assert qn.value == n.value and np.array_equal(q[:qn.value], out[:n.value])
