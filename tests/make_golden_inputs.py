"""Seeded inputs of the golden-vector cases and the two ways of evaluating them: through the CPU oracle (any machine)
and through the reference's own CUDA kernels (oracle/_ref/libperch_ref.so, GPU box).  tests/golden/make_golden.py
stores the latter; tests/test_oracle_cpu.py compares the former against the stored files.

The reference rasterizer is NOT deterministic on sm_70+ whenever two fragments land on one pixel (its spin lock has no
fence and the depth read may be served from a stale L1 line, image_renderer.cuh:146-159; the occlusion step races
with the z-test, :161-186).  Golden cases are therefore split:
  EXACT  frag_single_tri   every image holds ONE triangle => one fragment per pixel, no contention: pins the vertex
                           transform, viewport, bbox, inside test and depth rounding fragment by fragment
  EXACT  quad_occlusion    one quad (its two triangles meet only on the diagonal, with equal depths) over block-wise
                           varying source depth / labels / thresholds: pins the source-occlusion rule
  EXACT  tree_single_tri   the one-triangle images again, over block-wise varying source depth / labels, through the
                           build of the reference kernels with USE_TREE / USE_CLUTTER on (model.h:17-18): pins
                           pose_occluded, pose_occluded_other, the clutter ratio and the fragment count
  EXACT  cloud             compute_point_clouds in its three modes
  EXACT  cost_3dof/_6dof   reference compute_point_clouds + k=1 search + compute_costs fed with canonical images
  BUDGET render_3dof/_6dof full meshes over a real scene: the reference deviates on a few % of object pixels and from
                           run to run; the observed counts are stored as the budget
"""
import numpy as np

from perception_b200 import synth
from perception_b200.renderer import make_triangles, pose_to_mat4x4
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt

W, H = 160, 120
NAMES = ["frag_single_tri", "quad_occlusion", "render_3dof", "render_6dof", "cloud", "cost_3dof", "cost_6dof",
         "tree_single_tri"]
EXACT = {"frag_single_tri", "quad_occlusion", "cloud", "cost_3dof", "cost_6dof", "tree_single_tri"}


def _case():
    return make_case(600, 3, 42, W, H)


def _cost_inputs(orc, dof6):
    case = _case()
    s = case.scene
    stride = 2
    obs = observed_6dof(case, stride, orc) if dof6 else observed_3dof(case, stride, orc)
    poses, pmm, lab = poses_near_gt(case, 16)
    if dof6:
        tot = np.bincount(obs["label"], minlength=len(s.gt_poses)).astype(np.float32)[lab]
    else:
        tot = synth.cylinder_counts(s, obs["aos"], poses, pmm)
    return case, obs, poses, pmm, (lab if dof6 else None), tot, stride


def single_tri_inputs():
    """36 one-triangle 'models' (a box and a 24-triangle blob) x 3 poses."""
    case = _case()
    tris = np.concatenate([synth.make_box_mesh((0.09, 0.06, 0.11)), synth.make_mesh(24, seed=5, r0=0.12)])
    tris = tris.astype(tris.dtype)
    counts = np.ones(len(tris), np.int32)
    rng = np.random.default_rng(9)
    poses, pmm = [], []
    for k in range(3):
        T = np.eye(4)
        T[:3, :3] = synth.random_rotation(rng)
        T[:3, 3] = (rng.uniform(-0.1, 0.1), rng.uniform(-0.08, 0.08), rng.uniform(0.45, 0.8))
        for m in range(len(tris)):
            poses.append(pose_to_mat4x4(T))
            pmm.append(m)
    return case, tris, counts, np.stack(poses), np.array(pmm, np.int32)


def quad_inputs():
    case = _case()
    h = 0.07
    v = np.array([[[-h, -h, 0.60], [h, -h, 0.60], [h, h, 0.60]], [[-h, -h, 0.60], [h, h, 0.60], [-h, h, 0.60]]], np.float32)
    tris = make_triangles(v, np.array([[11, 22, 33], [44, 55, 66]]))
    src = np.zeros((H, W), np.int32)
    vals = [0, 58, 59, 60, 61, 1, 100, 57]
    for bx in range(W // 4):
        src[:, bx * 4:(bx + 1) * 4] = vals[bx % len(vals)]
    lab = np.zeros((H, W), np.uint8)
    for by in range(H // 6):
        lab[by * 6:(by + 1) * 6] = by % 3
    pose = pose_to_mat4x4(np.eye(4))[None]
    return case, tris, src, lab, pose


def tree_source():
    """Source depth in vertical stripes around the depths the one-triangle images reach (45-90 cm), labels in bands."""
    src = np.zeros((H, W), np.int32)
    vals = [0, 40, 52, 58, 61, 64, 70, 76, 90, 200, 1, 66]
    for bx in range(W // 3 + 1):
        src[:, bx * 3:(bx + 1) * 3] = vals[bx % len(vals)]
    lab = np.zeros((H, W), np.uint8)
    for by in range(H // 5):
        lab[by * 5:(by + 1) * 5] = by % 4
    return src, lab


def _cloud_modes(case):
    s = case.scene
    return [("plain", {}), ("label", {"label_mask": s.source_label}),
            ("bounds", {"bounds": synth.table_bounds(s), "cam_tf": np.linalg.inv(s.T_cam_table)})]


def _run(render, depth2cloud, cost_fn, orc, name, render_tree=None):
    """render(tris, counts, poses, pmm, src_depth, src_label, pose_label, thr) -> depth, rgb
    depth2cloud(**kw) -> dict; cost_fn(depth, rgb, pose_label, obs, tot, stride, dof6) -> dict
    render_tree(same as render) -> dict with depth, pose_occluded, pose_occluded_other, clutter_cost, total_points"""
    case = _case()
    s = case.scene
    if name == "tree_single_tri":
        case, tris, counts, poses, pmm = single_tri_inputs()
        src, lab = tree_source()
        out = {}
        for tag, sl, pl, thr in (("3dof", None, None, 1.0), ("3dof_thr5", None, None, 5.0),
                                 ("6dof", lab, (np.arange(len(poses)) % 3).astype(np.int32), 1.0)):
            o = render_tree(tris, counts, poses, pmm, src, sl, pl, thr)
            for k in ("depth", "pose_occluded", "pose_occluded_other", "clutter_cost", "total_points"):
                out[f"{tag}_{k}"] = o[k]
        return out
    if name == "frag_single_tri":
        case, tris, counts, poses, pmm = single_tri_inputs()
        d, c = render(tris, counts, poses, pmm, None, None, None, 1.0)
        return {"depth": d, "rgb": c}
    if name == "quad_occlusion":
        case, tris, src, lab, pose = quad_inputs()
        out = {}
        for i, thr in enumerate([0.0, 0.5, 1.0, 2.5]):
            d, c = render(tris, [2], pose, [0], src, None, None, thr)
            out[f"depth_thr{i}"], out[f"rgb_thr{i}"] = d, c
        for pl in (0, 1):
            d, c = render(tris, [2], pose, [0], src, lab, [pl], 1.0)
            out[f"depth_label{pl}"], out[f"rgb_label{pl}"] = d, c
        return out
    if name in ("render_3dof", "render_6dof"):
        poses, pmm, lab = poses_near_gt(case, 8)
        six = name == "render_6dof"
        d, c = render(s.tris, s.tris_model_count, poses, pmm, s.source_depth, s.source_label if six else None,
                      lab if six else None, 1.0)
        return {"depth": d, "rgb": c}
    if name == "cloud":
        out = {}
        for mode, kw in _cloud_modes(case):
            o = depth2cloud(**kw)
            out[f"{mode}_n"] = np.array(o["n"])
            out[f"{mode}_soa"] = o["soa"]
            out[f"{mode}_rgb"] = o["rgb"]
            if mode == "label":
                out["label_label"] = o["label"]
        return out
    dof6 = name == "cost_6dof"
    case, obs, poses, pmm, lab, tot, stride = _cost_inputs(orc, dof6)
    # canonical images (oracle) are the INPUT of the cloud+cost stage in both evaluations
    d, c = orc.render(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, s.source_depth,
                      s.source_label if dof6 else None, lab, 1.0)
    return cost_fn(d, c, lab, obs, tot, stride, dof6)


def run_oracle(orc, name):
    case = _case()
    s = case.scene

    def render(tris, counts, poses, pmm, src, src_label, pose_label, thr):
        return orc.render(tris, counts, poses, pmm, W, H, case.proj, src, src_label, pose_label, thr)

    def d2c(**kw):
        return orc.depth2cloud(s.source_depth, s.source_color, W, H, 1, case.cx, case.cy, case.fx, case.fy, 100.0, 2, **kw)

    def cost(d, c, lab, obs, tot, stride, dof6):
        r = orc.costs(d, c, W, H, case.cx, case.cy, case.fx, case.fy, 100.0, stride, lab, obs["aos"], obs["rgb"],
                      obs["label"] if dof6 else None, tot, 2 if dof6 else 0, True, 0.01 if dof6 else 0.0075, 20.0)
        return {"rendered_cost": r["rendered_cost"], "observed_cost": r["observed_cost"],
                "points_diff_cost": r["points_diff_cost"], "n_rend": np.array(int(r["counts"][0].sum()))}

    def render_tree(tris, counts, poses, pmm, src, src_label, pose_label, thr):
        o = orc.render_tree(tris, counts, poses, pmm, W, H, case.proj, src, src_label, pose_label, thr)
        # one fragment per pixel: the serial (per-fragment) and canonical (per-pixel) readings must coincide
        for k in ("pose_occluded", "pose_occluded_other", "clutter_points", "total_points"):
            assert np.array_equal(o["serial"][k], o["canon"][k]), k
        return {"depth": o["depth"], **o["serial"]}

    return _run(render, d2c, cost, orc, name, render_tree)


def run_reference(ref, orc, name):
    """Same cases through the reference CUDA kernels.  `orc` only builds INPUTS (observed cloud, canonical images)."""
    case = _case()
    s = case.scene

    def render(tris, counts, poses, pmm, src, src_label, pose_label, thr):
        return ref.render(tris, counts, poses, pmm, W, H, case.proj, src, None, src_label, pose_label, thr)

    def d2c(**kw):
        return ref.depth2cloud(s.source_depth, s.source_color, W, H, 1, case.cx, case.cy, case.fx, case.fy, 100.0, 2, **kw)

    def cost(d, c, lab, obs, tot, stride, dof6):
        q = ref.cost_from_images(d, c, W, H, case.cx, case.cy, case.fx, case.fy, 100, stride, lab, obs["aos"], obs["rgb"],
                                 obs["label"] if dof6 else None, tot, 2 if dof6 else 0, True, 0.01 if dof6 else 0.0075, 20.0)
        return {"rendered_cost": q["rendered_cost"], "observed_cost": q["observed_cost"],
                "points_diff_cost": q["points_diff_cost"], "n_rend": np.array(q["n_rend"])}

    def render_tree(tris, counts, poses, pmm, src, src_label, pose_label, thr):
        from oracle_lib import Reference
        return Reference(tree=True).render_tree(tris, counts, poses, pmm, W, H, case.proj, src, src_label, pose_label, thr)

    return _run(render, d2c, cost, orc, name, render_tree)
