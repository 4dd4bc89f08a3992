"""The BASELINE.json workloads exactly as bench.py builds and times them (bench.build_workload: C1 3-DoF 5k-triangle,
C2 16 384 6-DoF poses x 20k triangles, C3 5 meshes, C5 21 meshes), at FULL size, held against

  (a) the CPU oracle (oracle_score_poses = oracle_render + oracle_costs per pose, following image_renderer.cuh:59-210,
      compute_point_clouds.cuh:37-184, compute_costs.cuh:161-290) on a >= 256-pose sample spread over the batch, and
  (b) the library's unfused kernels (perch_render_cost_unfused: every pixel rasterised without any culling, clouds,
      brute-force nearest neighbour) on EVERY pose of the batch.

This is the regime in which the fused kernel's three approximate culling levels decide what reaches the exact path."""
import numpy as np
import pytest

import bench
from conftest import has_gpu
from oracle_lib import Oracle
from perception_b200 import _capi, synth
from perception_b200.renderer import Renderer

pytestmark = pytest.mark.gpu

N_ORACLE = 512


def _kw(wl):
    return dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"],
                color_distance_threshold=20.0, occlusion_threshold=1.0)


def _oracle_sample(wl, sel):
    s = wl["scene"]
    lab = wl["plabel"][sel] if wl["plabel"] is not None else None
    return Oracle().score_poses(s.tris, s.tris_model_count, wl["poses"][sel], wl["pmm"][sel], bench.W, bench.H, wl["proj"],
                                s.source_depth, s.source_label if wl["dof6"] else None, lab, synth.CX, synth.CY, synth.FX,
                                synth.FY, synth.DEPTH_FACTOR, wl["stride"], wl["obs"]["result_cloud_eigen"],
                                wl["obs"]["result_cloud_color"], wl["obs_label"], wl["tot"][sel], wl["cost_type"], True,
                                wl["res"], 20.0)


def _same(a, b, idx=None):
    for k in ("counts", "rendered_cost", "observed_cost", "points_diff_cost"):
        x = a[k] if idx is None else a[k][..., idx]
        y = b[k]
        assert x.shape == y.shape, k
        bad = np.nonzero(~((x == y) | (np.isnan(x) & np.isnan(y))))
        assert len(bad[0]) == 0, f"{k}: {len(bad[0])} mismatches, first at {tuple(int(v[0]) for v in bad)}"


@pytest.mark.parametrize("name", ["c1", "c2", "c3", "c5"])
def test_bench_workload_vs_oracle_and_unfused(name):
    r, wl = bench.build_workload(name, 0, 0, None)
    try:
        P = wl["P"]
        fused = r.render_cuda_multi_unified("COST", wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **_kw(wl))
        assert (fused["counts"][0] > 0).mean() > 0.5, "most poses of the workload must render something"
        # (a) oracle on a sample spread over the whole batch (all meshes of C3 / C5 are hit)
        sel = np.unique(np.linspace(0, P - 1, min(N_ORACLE, P)).astype(np.int64))
        assert len(sel) >= min(256, P)
        _same(fused, _oracle_sample(wl, sel), sel)
        # (b) every pose against the unfused device path
        unf = r.render_cost_unfused(wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], **_kw(wl))
        _same(fused, unf)
    finally:
        r.close()


@pytest.mark.parametrize("name", ["c2", "c5"])
def test_scene_call_topk_at_full_size(name):
    """perch_score_scene on the bench batches: C2's 16 384 poses are exactly what the select kernel keeps in registers
    (every slot used), C5's 125 000 poses x 21 models take its streaming path; both must return what the caller's rule
    (perch_topk, search_env.cpp:2542-2583) returns on the same costs."""
    from perception_b200.renderer import topk
    r, wl = bench.build_workload(name, 0, 0, None)
    try:
        n_models = len(wl["scene"].tris_model_count)
        hits, costs = r.score_scene(wl["poses"], wl["pmm"], wl["plabel"], wl["tot"], n_models=n_models, k=8, **_kw(wl))
        want = topk(costs["rendered_cost"], costs["observed_cost"], wl["pmm"], n_models, k=8)
        assert np.array_equal(hits, want)
        # (some models of C5 have no pose that passes the caller's |rendered - observed| < 30 gate: their rows are all pads)
        assert (want["cost"][:, 0] < np.iinfo(np.int32).max).mean() > 0.5, "most models of the workload have a valid best pose"
    finally:
        r.close()


def test_pose_labels_without_observed_labels_is_an_error_not_a_fault():
    """6-DoF call (pose labels) after set_observed(label=None): PERCH_ERR_INVALID, and the context stays usable."""
    from scenes import make_case, observed_3dof, poses_near_gt
    orc = Oracle()
    case = make_case(600, 3, 42)
    s = case.scene
    obs = observed_3dof(case, 4, orc)
    poses, pmm, lab = poses_near_gt(case, 8)
    tot = np.ones(8, np.float32)
    with Renderer(0) as r:
        r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
        r.set_models(s.tris, s.tris_model_count)
        r.set_scene(s.source_depth, s.source_color, s.source_label)
        r.set_observed(obs["aos"], obs["rgb"], None)
        for call in (lambda: r.render_cuda_multi_unified("COST", poses, pmm, lab, tot, stride=4, cost_type=2),
                     lambda: r.gicp(poses, pmm, lab, stride=4),
                     lambda: r.render_cost_unfused(poses, pmm, lab, tot, stride=4, cost_type=2)):
            with pytest.raises(_capi.PerchError) as e:
                call()
            assert e.value.code == _capi.PERCH_ERR_INVALID
        ok = r.render_cuda_multi_unified("COST", poses, pmm, None, tot, stride=4, cost_type=0)  # no sticky CUDA error
        assert (ok["counts"][0] > 0).any()
