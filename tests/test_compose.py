"""Composition mode through the C-ABI: perch_compose_source, perch_tree_flags.

Parity: the composed source image and every cost computed against it are bit-exact vs the CPU oracle; the tree /
clutter vectors are bit-exact vs the oracle's canonical (per covered pixel) reading on full meshes, and vs the
reference's own kernels built with USE_TREE / USE_CLUTTER on (oracle/_ref/libperch_ref_tree.so, golden
tests/golden/tree_single_tri.npz) where every pixel receives one fragment and the reference is deterministic."""
from pathlib import Path

import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import REF_TREE_SO, Oracle, Reference
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt

pytestmark = pytest.mark.gpu

if has_gpu():
    from perception_b200 import synth
    from perception_b200.renderer import Renderer

GOLDEN = Path(__file__).resolve().parent / "golden"
TREE_KEYS = ("pose_occluded", "pose_occluded_other", "clutter_cost", "total_points")


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _renderer(case, with_label=False):
    s = case.scene
    r = Renderer(0)
    r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
    r.set_models(s.tris, s.tris_model_count)
    r.set_scene(s.source_depth, s.source_color, s.source_label if with_label else None)
    return r


def _same(a, b):
    return np.array_equal(a, b, equal_nan=True)


def test_tree_flags_one_fragment_per_pixel(orc):
    """product == oracle (both readings) == reference kernels with USE_TREE / USE_CLUTTER on == committed golden."""
    import make_golden_inputs as gi
    case, tris, counts, poses, pmm = gi.single_tri_inputs()
    src, lab = gi.tree_source()
    golden = np.load(GOLDEN / "tree_single_tri.npz") if (GOLDEN / "tree_single_tri.npz").exists() else None
    ref = Reference(tree=True) if REF_TREE_SO.exists() else None
    with Renderer(0) as r:
        r.set_camera(gi.W, gi.H, case.proj, case.cx, case.cy, case.fx, case.fy, 100.0)
        r.set_models(tris, counts)
        for tag, sl, pl, thr in (("3dof", None, None, 1.0), ("3dof_thr5", None, None, 5.0),
                                 ("6dof", lab, (np.arange(len(poses)) % 3).astype(np.int32), 1.0)):
            r.set_scene(src, None, sl)
            out = r.tree_flags(poses, pmm, pl, thr)
            o = orc.render_tree(tris, counts, poses, pmm, gi.W, gi.H, case.proj, src, sl, pl, thr)
            assert o["canon"]["pose_occluded"].sum() > 5 and o["canon"]["pose_occluded_other"].sum() > 5
            assert (o["canon"]["clutter_points"] > 0).sum() > 5
            for k in TREE_KEYS:
                assert _same(out[k], o["canon"][k]), (tag, k)
                assert _same(out[k], o["serial"][k]), (tag, k)
                if golden is not None:
                    assert _same(out[k], golden[f"ref_{tag}_{k}"]), (tag, k, "golden")
            if ref is not None:
                q = ref.render_tree(tris, counts, poses, pmm, gi.W, gi.H, case.proj, src, sl, pl, thr)
                for k in TREE_KEYS:
                    assert _same(out[k], q[k]), (tag, k, "live reference")
                assert np.array_equal(q["depth"], o["depth"])


@pytest.mark.parametrize("dof6", [False, True])
def test_tree_flags_full_meshes(orc, dof6):
    """Many fragments per pixel: the product equals the canonical reading bit for bit.  The per-fragment reading in
    triangle order (one legal execution of the reference) must agree on pose_occluded (order independent) and can only
    have more fragments than covered pixels."""
    case = make_case(2000, 4, 7)
    s = case.scene
    poses, pmm, lab = poses_near_gt(case, 24, trans_sigma=0.03)
    pl = lab if dof6 else None
    with _renderer(case, with_label=dof6) as r:
        out = r.tree_flags(poses, pmm, pl, 1.0)
    o = orc.render_tree(s.tris, s.tris_model_count, poses, pmm, case.width, case.height, case.proj, s.source_depth,
                        s.source_label if dof6 else None, pl, 1.0)
    for k in TREE_KEYS:
        assert _same(out[k], o["canon"][k]), k
    assert np.array_equal(out["pose_occluded"], o["serial"]["pose_occluded"])
    assert (o["serial"]["total_points"] >= out["total_points"]).all()
    assert (o["serial"]["pose_occluded_other"] >= out["pose_occluded_other"]).all()
    assert np.median(out["total_points"]) > 100 and out["pose_occluded_other"].sum() > 0 and out["pose_occluded"].sum() > 0


def _compose_expected(orc, case, poses, pmm, plabel, mode, with_label):
    """The reference's way: render the placed objects into ONE image against an empty source (PrintStateGPU,
    search_env.cpp:1718-1781), then merge on the host."""
    s = case.scene
    W, H = case.width, case.height
    placed = np.zeros((H, W), np.int32)
    owner = np.full((H, W), -1, np.int64)
    for i in range(len(poses)):  # one image per pose so that the owner of every pixel is known
        d, _ = orc.render(s.tris, s.tris_model_count, poses[i:i + 1], pmm[i:i + 1], W, H, case.proj)
        d = d[0]
        win = (d > 0) & ((placed == 0) | (d < placed))
        placed[win] = d[win]
        owner[win] = i
    plab = np.zeros((H, W), np.uint8)
    if plabel is not None:
        plab[owner >= 0] = (np.asarray(plabel)[owner[owner >= 0]] + 1).astype(np.uint8)
    if mode == "replace":
        return placed, plab
    base = s.source_depth.reshape(H, W)
    base_lab = s.source_label.reshape(H, W) if with_label else np.zeros((H, W), np.uint8)
    take = (placed > 0) & ~((base > 0) & (base <= placed))
    return np.where(take, placed, base), np.where(take, plab, base_lab).astype(np.uint8)


@pytest.mark.parametrize("mode,dof6", [("replace", False), ("merge", False), ("merge", True), ("replace", True)])
def test_compose_source_and_costs_against_it(orc, mode, dof6):
    case = make_case(2000, 4, 7)
    s = case.scene
    W, H = case.width, case.height
    stride = 4
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    cand, cmm, clab = poses_near_gt(case, 24, trans_sigma=0.02)
    # placed objects: two instances pushed towards the camera so that they hide parts of the candidates
    placed, pmm_p, plab = poses_near_gt(case, 2)
    placed = placed.copy()
    placed[:, 2, 3] -= 6.0  # cm
    if dof6:
        tot = np.bincount(obs["label"], minlength=len(s.gt_poses)).astype(np.float32)[clab]
    else:
        tot = synth.cylinder_counts(s, obs["aos"], cand, cmm)
    kw = dict(stride=stride, cost_type=2 if dof6 else 0, sensor_resolution=0.01 if dof6 else 0.0075, occlusion_threshold=1.0)
    with _renderer(case, with_label=dof6) as r:
        r.set_observed(obs["aos"], obs["rgb"], obs["label"] if dof6 else None)
        before = r.render_cuda_multi_unified("COST", cand, cmm, clab if dof6 else None, tot, **kw)
        r.compose_source(placed, pmm_p, plab if dof6 else None, mode=mode)
        got_d, got_l = r.get_source_depth(), r.get_source_label()
        after = r.render_cuda_multi_unified("COST", cand, cmm, clab if dof6 else None, tot, **kw)
        img = r.render_cuda_multi_unified("RENDER", cand[:6], cmm[:6], clab[:6] if dof6 else None)
        r.compose_source(mode="none")
        restored = r.render_cuda_multi_unified("COST", cand, cmm, clab if dof6 else None, tot, **kw)
        assert np.array_equal(r.get_source_depth().reshape(-1), s.source_depth.reshape(-1))
    exp_d, exp_l = _compose_expected(orc, case, placed, pmm_p, plab if dof6 else None, mode, dof6)
    assert (exp_d != s.source_depth.reshape(H, W)).sum() > 500
    assert np.array_equal(got_d, exp_d)
    if dof6:
        assert got_l is not None and np.array_equal(got_l, exp_l)
    # every stage now tests occlusion against the composed image: same results as the oracle fed with it as source
    src_l = exp_l if dof6 else None
    d_o, c_o = orc.render(s.tris, s.tris_model_count, cand, cmm, W, H, case.proj, exp_d, src_l, clab if dof6 else None, 1.0)
    o = orc.costs(d_o, c_o, W, H, case.cx, case.cy, case.fx, case.fy, 100.0, stride, clab if dof6 else None, obs["aos"],
                  obs["rgb"], obs["label"] if dof6 else None, tot, kw["cost_type"], True, kw["sensor_resolution"], 20.0)
    assert np.array_equal(after["counts"], o["counts"])
    for k in ("rendered_cost", "observed_cost", "points_diff_cost"):
        assert _same(after[k], o[k]), k
    assert np.array_equal(img["result_depth"], d_o[:6]) and np.array_equal(img["result_color"], c_o[:, :6])
    assert not np.array_equal(before["counts"], after["counts"])  # the placed objects do hide candidates
    assert np.array_equal(before["counts"], restored["counts"])


def test_greedy_multi_object_search_with_composition(orc):
    """BASELINE configs[2] in miniature: objects are committed one at a time; each next object is scored with the
    committed ones rendered into the source (merge mode).  Checks the whole sequence against the oracle."""
    case = make_case(600, 3, 42)
    s = case.scene
    W, H = case.width, case.height
    stride = 4
    obs = observed_3dof(case, stride, orc)
    committed, committed_mm = [], []
    src = s.source_depth.reshape(H, W).copy()
    with _renderer(case) as r:
        r.set_observed(obs["aos"], obs["rgb"], None)
        for g in range(len(s.gt_poses)):
            cand, cmm, lab = poses_near_gt(case, 30, seed=20 + g, trans_sigma=0.015)
            sel = lab == g
            cand, cmm = cand[sel], cmm[sel]
            tot = synth.cylinder_counts(s, obs["aos"], cand, cmm)
            out = r.render_cuda_multi_unified("COST", cand, cmm, None, tot, stride=stride, sensor_resolution=0.0075)
            d_o, c_o = orc.render(s.tris, s.tris_model_count, cand, cmm, W, H, case.proj, src, None, None, 1.0)
            o = orc.costs(d_o, c_o, W, H, case.cx, case.cy, case.fx, case.fy, 100.0, stride, None, obs["aos"], obs["rgb"],
                          None, tot, 0, True, 0.0075, 20.0)
            assert np.array_equal(out["counts"], o["counts"]), g
            best = int(np.nanargmin(np.where(out["rendered_cost"] < 0, np.inf, out["rendered_cost"] + out["observed_cost"])))
            committed.append(cand[best]); committed_mm.append(cmm[best])
            r.compose_source(np.stack(committed), np.array(committed_mm, np.int32), mode="merge")
            src, _ = _compose_expected(orc, case, np.stack(committed), np.array(committed_mm, np.int32), None, "merge", False)
            assert np.array_equal(r.get_source_depth(), src), g
