"""Full-resolution rasteriser (perception_b200/csrc/raster_full.cu) against the per-pose kernel it replaces for stages
RENDER / DEBUG, perch_compose_source and perch_tree_flags: both implement the canonical raster semantics, so every
output must agree bit for bit -- across pose chunks, with every triangle forced through the deferred (warp per triangle)
path, and with none.  Parity with the oracle / reference goldens is covered by test_parity_gpu.py and test_compose.py,
which run on the new path by default."""
import os

import numpy as np
import pytest

from conftest import has_gpu
from scenes import make_case, poses_near_gt

pytestmark = pytest.mark.gpu

if has_gpu():
    from perception_b200 import synth
    from perception_b200.renderer import Renderer

IMG_KEYS = ("result_depth", "result_color")
TREE_KEYS = ("pose_occluded", "pose_occluded_other", "clutter_cost", "total_points")


class _env:
    def __init__(self, **kv):
        self.kv = {k: str(v) for k, v in kv.items()}

    def __enter__(self):
        self.old = {k: os.environ.get(k) for k in self.kv}
        os.environ.update(self.kv)

    def __exit__(self, *exc):
        for k, v in self.old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _renderer(case, with_label):
    s = case.scene
    r = Renderer(0)
    r.set_camera(case.width, case.height, case.proj, case.cx, case.cy, case.fx, case.fy, synth.DEPTH_FACTOR)
    r.set_models(s.tris, s.tris_model_count)
    r.set_scene(s.source_depth, s.source_color, s.source_label if with_label else None)
    return r


VARIANTS = ({}, {"PERCH_FULL_CHUNK_MB": 5}, {"PERCH_FULL_BIG_CELLS": 1}, {"PERCH_FULL_BIG_CELLS": 1 << 30},
            {"PERCH_FULL_CHUNK_MB": 1, "PERCH_FULL_BIG_CELLS": 9}, {"PERCH_FULL_LANE_CELLS": 0}, {"PERCH_FULL_CHUNK_MB": 3},
            {"PERCH_FULL_LANE_CELLS": 1 << 30, "PERCH_FULL_BIG_CELLS": 1 << 30})


@pytest.mark.parametrize("dof6", [False, True])
def test_render_images_equal_per_pose_kernel(dof6):
    case = make_case(3000, 4, 11)
    poses, pmm, lab = poses_near_gt(case, 37, trans_sigma=0.04)
    # one pose far off-screen, one straddling the image border
    poses[3][0, 3] += 500.0
    poses[5][0, 3] += 25.0
    with _renderer(case, dof6) as r:
        pl = lab if dof6 else None
        with _env(PERCH_IMAGE_LEGACY=1):
            want = r.render_cuda_multi_unified("RENDER", poses, pmm, pl)
            want1 = r.render_cuda_multi_unified("RENDER", poses, pmm, pl, single_result_image=True)
        assert (want["result_depth"] > 0).sum() > 1000
        for v in VARIANTS:
            with _env(**v):
                got = r.render_cuda_multi_unified("RENDER", poses, pmm, pl)
                got1 = r.render_cuda_multi_unified("RENDER", poses, pmm, pl, single_result_image=True)
            for k in IMG_KEYS:
                assert np.array_equal(got[k], want[k]), (v, k)
                assert np.array_equal(got1[k], want1[k]), (v, k, "single")
        # repeated calls leave the key images clean
        again = r.render_cuda_multi_unified("RENDER", poses[:5], pmm[:5], pl[:5] if dof6 else None)
        for k in IMG_KEYS:
            assert np.array_equal(again[k], want[k][:5] if k == "result_depth" else want[k][:, :5]), k


@pytest.mark.parametrize("dof6", [False, True])
def test_tree_flags_and_compose_equal_per_pose_kernel(dof6):
    case = make_case(2500, 4, 5)
    poses, pmm, lab = poses_near_gt(case, 29, trans_sigma=0.03)
    pl = lab if dof6 else None
    with _renderer(case, dof6) as r:
        with _env(PERCH_IMAGE_LEGACY=1):
            want = r.tree_flags(poses, pmm, pl, 1.0)
            r.compose_source(poses[:3], pmm[:3], pl[:3] if dof6 else None, "merge")
            want_src, want_lab = r.get_source_depth(), r.get_source_label()
            r.compose_source(None, None, None, "none")
        assert want["total_points"].max() > 100
        for v in VARIANTS:
            with _env(**v):
                got = r.tree_flags(poses, pmm, pl, 1.0)
                r.compose_source(poses[:3], pmm[:3], pl[:3] if dof6 else None, "merge")
                src, lb = r.get_source_depth(), r.get_source_label()
                r.compose_source(None, None, None, "none")
            for k in TREE_KEYS:
                assert np.array_equal(got[k], want[k], equal_nan=True), (v, k)
            assert np.array_equal(src, want_src), v
            assert (lb is None) == (want_lab is None) and (lb is None or np.array_equal(lb, want_lab)), v


def test_debug_stage_images_and_costs_unchanged():
    from scenes import observed_3dof
    case = make_case(2000, 3, 2)
    poses, pmm, _ = poses_near_gt(case, 12, trans_sigma=0.02)
    with _renderer(case, False) as r:
        obs = observed_3dof(case, 5)
        r.set_observed(obs["aos"], obs["rgb"], None)
        with _env(PERCH_IMAGE_LEGACY=1):
            want = r.render_cuda_multi_unified("DEBUG", poses, pmm, None, stride=5)
        got = r.render_cuda_multi_unified("DEBUG", poses, pmm, None, stride=5)
        for k in want:
            if isinstance(want[k], np.ndarray):
                assert np.array_equal(got[k], want[k], equal_nan=True), k
