"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/perch_b200.h
declares, its host-only entry points behave like the reference's host code, and it refuses to run without a GPU."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from conftest import has_gpu
from oracle_lib import Oracle
from perception_b200 import _capi, synth
from perception_b200.renderer import TRIANGLE_DTYPE, compute_proj, topk

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "perch_b200.h").read_text()


def declared_symbols():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    return sorted(set(re.findall(r"\b(perch_[a-z0-9_]+)\s*\(", body)))


def test_library_exports_every_declared_symbol():
    lib = _capi.load_library()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/perch_b200.h but not exported"
        assert n in _capi.SIGNATURES, f"{n} has no ctypes signature"
    assert lib.perch_abi_version() == int(re.search(r"PERCH_B200_ABI_VERSION (\d+)", HEADER).group(1))


def test_struct_layouts_match_reference():
    # cuda_renderer::Model::Triangle is 64 bytes with the colour at offset 40 (model.h:62-73, size_t alignment)
    assert TRIANGLE_DTYPE.itemsize == 64 and TRIANGLE_DTYPE.fields["color"][1] == 40
    assert C.sizeof(_capi.perch_hit) == 16
    assert C.sizeof(_capi.perch_params) == 48


def test_compute_proj_equals_oracle():
    for K, w, h in [(synth.camera_K(), 640, 480), (np.array([[300.5, 0.25, 161.2], [0, 299.1, 118.7], [0, 0, 1]], np.float32), 320, 240)]:
        a = compute_proj(K, w, h, 10, 10000)
        b = Oracle().compute_proj(K, w, h, 10, 10000)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_topk_follows_caller_rule():
    # search_env.cpp:2016-2036 + 2542-2583: cost=(int)(r+o); skip (int)r<0; |int(r)-int(o)| < 30; strict < => first wins
    r = np.array([10.9, 10.2, -1.0, 50.0, 5.0, 3.5, 3.9], np.float32)
    o = np.array([10.9, 10.1, 5.0, 10.0, 60.0, 3.0, 3.0], np.float32)
    m = np.array([0, 0, 0, 0, 0, 1, 1], np.int32)
    t = topk(r, o, m, n_models=3, k=2, pose_id_offset=100)
    assert t["cost"][0].tolist() == [20, 21] and t["pose_id"][0].tolist() == [101, 100]
    assert t["cost"][1].tolist() == [6, 6] and t["pose_id"][1].tolist() == [105, 106]  # tie -> lowest pose id first
    assert t["cost"][2].tolist() == [2**31 - 1] * 2 and t["pose_id"][2].tolist() == [-1, -1]
    nan = topk(np.array([1, np.nan], np.float32), np.array([np.nan, 1], np.float32), np.zeros(2, np.int32), 1, 1)
    assert nan["pose_id"][0, 0] == -1


def test_mesh_preprocessing_keeps_every_triangle_bit_for_bit():
    """perch_set_models re-files the triangle soup as quads (edge-sharing pairs) in clusters: every triangle must be
    stored exactly once, with its coordinates and vertex ORDER intact (the exact path depends on the order), inside its
    cluster's bounding sphere -- for closed meshes, soups without shared vertices, degenerate and non-finite input."""
    lib = _capi.load_library()
    rng = np.random.default_rng(3)
    blob, box = synth.make_mesh(2000, seed=1000, r0=0.08), synth.make_box_mesh((0.05, 0.035, 0.06))
    soup = synth.make_mesh(300, seed=7, r0=0.05).copy()
    soup["v"] += rng.normal(0, 1e-6, soup["v"].shape).astype(np.float32)      # no two vertices identical any more
    odd = synth.make_mesh(60, seed=9, r0=0.05).copy()
    odd["v"][3, 1] = odd["v"][3, 0]                                            # zero-area triangle with a repeated vertex
    odd["v"][7] = odd["v"][6]                                                  # exact duplicate triangle
    odd["v"][11, 2, 0] = np.nan
    odd["v"][12, 0, 1] = np.inf
    for models in ([blob], [box], [soup], [odd], [blob, box, odd, soup], [blob[:1]], [blob, blob[:0], box]):
        tris = np.ascontiguousarray(np.concatenate(models), dtype=TRIANGLE_DTYPE)
        counts = np.array([len(m) for m in models], np.int32)
        out = np.zeros(6, np.int64)
        rc = lib.perch_cluster_stats(tris.ctypes.data_as(C.c_void_p), len(tris), counts.ctypes.data_as(C.c_void_p), len(counts),
                                     out.ctypes.data_as(C.c_void_p))
        assert rc == 0
        n_cl, paired, single, mismatch, missing, outside = out.tolist()
        assert paired + single == len(tris) and mismatch == 0 and missing == 0 and outside == 0, out
        assert n_cl * 4 >= len(tris)
    # a closed manifold mesh pairs almost completely, a soup not at all
    out = np.zeros(6, np.int64)
    t = np.ascontiguousarray(blob, dtype=TRIANGLE_DTYPE); c1 = np.array([len(t)], np.int32)
    lib.perch_cluster_stats(t.ctypes.data_as(C.c_void_p), len(t), c1.ctypes.data_as(C.c_void_p), 1, out.ctypes.data_as(C.c_void_p))
    assert out[1] >= 0.9 * len(t), out
    t = np.ascontiguousarray(soup, dtype=TRIANGLE_DTYPE); c1 = np.array([len(t)], np.int32)
    lib.perch_cluster_stats(t.ctypes.data_as(C.c_void_p), len(t), c1.ctypes.data_as(C.c_void_p), 1, out.ctypes.data_as(C.c_void_p))
    assert out[1] == 0 and out[2] == len(t), out


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU behaviour")
def test_create_fails_loudly_without_gpu():
    lib = _capi.load_library()
    h = C.c_void_p()
    rc = lib.perch_create(0, C.byref(h))
    assert rc == _capi.PERCH_ERR_NO_DEVICE and not h.value
    assert b"no CPU fallback" in lib.perch_last_error()
    from perception_b200.renderer import Renderer
    with pytest.raises(_capi.PerchError):
        Renderer(0)


def test_missing_library_raises(tmp_path):
    with pytest.raises(ImportError):
        _capi.load_library(tmp_path / "libperch_b200.so")


def test_pose_list_intake_round_trip(tmp_path):
    """poses.txt of the 6-DoF flow (search_env.cpp:7109-7130): `x y z qx qy qz qw` per line -> perch_state records."""
    from perception_b200 import states as S
    rng = np.random.default_rng(2)
    rows = rng.normal(0, 1, (50, 7))
    rows[:, 3:] /= np.linalg.norm(rows[:, 3:], axis=1, keepdims=True)
    text = "\n".join(" ".join(repr(float(v)) for v in r) for r in rows) + "\n"
    p = tmp_path / "poses.txt"
    p.write_text(text)
    got = S.load_pose_list(str(p), 3, 2)
    want = S.states_from_pose_lines(3, rows, 2)
    assert got.tobytes() == want.tobytes()          # std::stod round-trips repr() exactly
    # windows line ends, no trailing newline, extra columns (the reference reads the first seven), scientific notation
    got2 = S.load_pose_list("1 2 3 0 0 0 1 99\r\n1e-1 -2.5E0 +3 0.5 0.5 0.5 0.5", 0, -1)
    assert len(got2) == 2 and got2["x"].tolist() == [1.0, 0.1] and got2["y"][1] == -2.5 and got2["qw"].tolist() == [1.0, 0.5]
    assert len(S.load_pose_list("", 0, -1)) == 0
    for bad in ("1 2 3 4 5 6\n", "1 2 3 0 0  0 1\n", "1 2 x 0 0 0 1\n"):
        with pytest.raises(_capi.PerchError) as e:
            S.load_pose_list(bad, 0, -1)
        assert e.value.code == _capi.PERCH_ERR_INVALID and "line 1" in str(e.value)
