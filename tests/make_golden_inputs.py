"""Seeded inputs of the golden-vector cases and the two ways of evaluating them: through the CPU oracle (any machine)
and through the reference's own CUDA kernels (oracle/_ref/libperch_ref.so, GPU box).  tests/golden/make_golden.py
stores the latter; tests/test_oracle_cpu.py compares the former against the stored files."""
import numpy as np

from perception_b200 import synth
from scenes import make_case, observed_3dof, observed_6dof, poses_near_gt

W, H = 160, 120
NAMES = ["render_3dof", "render_6dof", "cloud", "cost_3dof", "cost_6dof"]


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


def run_oracle(orc, name):
    case = _case()
    s = case.scene
    if name in ("render_3dof", "render_6dof"):
        poses, pmm, lab = poses_near_gt(case, 8)
        six = name == "render_6dof"
        d, c = orc.render(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, s.source_depth,
                          s.source_label if six else None, lab if six else None, 1.0)
        return {"depth": d, "rgb": c}
    if name == "cloud":
        out = {}
        for mode, kw in _cloud_modes(case):
            o = orc.depth2cloud(s.source_depth, s.source_color, W, H, 1, case.cx, case.cy, case.fx, case.fy, 100.0, 2, **kw)
            out[f"{mode}_n"] = np.array(o["n"])
            out[f"{mode}_soa"] = o["soa"]
            out[f"{mode}_rgb"] = o["rgb"]
            if mode == "label":
                out["label_label"] = o["label"]
        return out
    dof6 = name == "cost_6dof"
    case, obs, poses, pmm, lab, tot, stride = _cost_inputs(orc, dof6)
    d, c = orc.render(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, s.source_depth,
                      s.source_label if dof6 else None, lab, 1.0)
    r = orc.costs(d, c, W, H, case.cx, case.cy, case.fx, case.fy, 100.0, stride, lab, obs["aos"], obs["rgb"],
                  obs["label"] if dof6 else None, tot, 2 if dof6 else 0, True, 0.01 if dof6 else 0.0075, 20.0)
    return {"rendered_cost": r["rendered_cost"], "observed_cost": r["observed_cost"],
            "points_diff_cost": r["points_diff_cost"], "n_rend": np.array(int(r["counts"][0].sum()))}


def _cloud_modes(case):
    s = case.scene
    return [("plain", {}), ("label", {"label_mask": s.source_label}),
            ("bounds", {"bounds": synth.table_bounds(s), "cam_tf": np.linalg.inv(s.T_cam_table)})]


def run_reference(ref, orc, name):
    """Same cases through the reference CUDA kernels.  `orc` is only used to build the observed-cloud INPUTS."""
    case = _case()
    s = case.scene
    if name in ("render_3dof", "render_6dof"):
        poses, pmm, lab = poses_near_gt(case, 8)
        six = name == "render_6dof"
        d, c = ref.render(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, s.source_depth, s.source_color,
                          s.source_label if six else None, lab if six else None, 1.0)
        return {"depth": d, "rgb": c}
    if name == "cloud":
        out = {}
        for mode, kw in _cloud_modes(case):
            o = ref.depth2cloud(s.source_depth, s.source_color, W, H, 1, case.cx, case.cy, case.fx, case.fy, 100.0, 2, **kw)
            out[f"{mode}_n"] = np.array(o["n"])
            out[f"{mode}_soa"] = o["soa"]
            out[f"{mode}_rgb"] = o["rgb"]
            if mode == "label":
                out["label_label"] = o["label"]
        return out
    dof6 = name == "cost_6dof"
    case, obs, poses, pmm, lab, tot, stride = _cost_inputs(orc, dof6)
    q = ref.render_cost(s.tris, s.tris_model_count, poses, pmm, W, H, case.proj, s.source_depth, s.source_color,
                        s.source_label if dof6 else None, lab, stride, 100, case.cx, case.cy, case.fx, case.fy, obs["aos"],
                        obs["rgb"], obs["label"] if dof6 else None, tot, 2 if dof6 else 0, True, 0.01 if dof6 else 0.0075,
                        20.0, 1.0)
    return {"rendered_cost": q["rendered_cost"], "observed_cost": q["observed_cost"],
            "points_diff_cost": q["points_diff_cost"], "n_rend": np.array(q["n_rend"])}
