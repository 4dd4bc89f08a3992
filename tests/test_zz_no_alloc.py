"""perch_depth2cloud_static -- the allocation-free variant of the depth image -> cloud kernels (scratch in a file-scope
device array, no cudaMalloc in the call) -- against the existing perch_depth2cloud entry point, through the C-ABI.

Collected last on purpose: a failure here must not hide the parity tests before it.  The calls
run back to back on one context with different shapes, so the second and third call see the previous call's contents in
the shared static array."""
import ctypes as C

import numpy as np
import pytest

from perception_b200 import _capi
from perception_b200.renderer import Renderer
from scenes import make_case

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 20  # PERCH_STATIC_CLOUD_MAX_ROWS


class _Dev:
    """Device buffers through the library's own helpers (no torch, no other CUDA binding)."""

    def __init__(self, lib, h):
        self.lib, self.h, self.ptrs = lib, h, []

    def alloc(self, nbytes):
        p = C.c_void_p()
        _capi.check(self.lib, self.lib.perch_device_alloc(self.h, max(int(nbytes), 1), C.byref(p)))
        self.ptrs.append(p)
        return p

    def put(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.alloc(arr.nbytes)
        _capi.check(self.lib, self.lib.perch_memcpy_h2d(self.h, p, arr.ctypes.data_as(C.c_void_p), arr.nbytes))
        return p

    def get(self, p, shape, dtype):
        out = np.empty(shape, dtype)
        if out.nbytes:
            _capi.check(self.lib, self.lib.perch_memcpy_d2h(self.h, out.ctypes.data_as(C.c_void_p), p, out.nbytes))
        return out

    def free(self):
        for p in self.ptrs:
            self.lib.perch_device_free(self.h, p)
        self.ptrs = []


def _static_call(r, dev, depth, label, W, H, n_img, cam, stride, capacity):
    lib, h = r._lib, r._h
    d_depth = dev.put(depth.astype(np.int32))
    d_label = dev.put(label.astype(np.uint8)) if label is not None else None
    d_aos, d_soa = dev.alloc(capacity * 12), dev.alloc(capacity * 12)
    d_map, d_lab = dev.alloc(capacity * 4), dev.alloc(capacity * 4)
    n = C.c_int(-1)
    rc = lib.perch_depth2cloud_static(h, d_depth, d_label, W, H, n_img, *cam, 100.0, stride, capacity, d_aos, d_soa, d_map,
                                      d_lab if label is not None else None, C.byref(n))
    out = {"rc": rc, "n": n.value}
    if rc == 0:
        k = n.value
        out["aos"] = dev.get(d_aos, (capacity, 3), np.float32)[:k]
        out["soa"] = dev.get(d_soa, (3, capacity), np.float32)[:, :k]
        out["map"] = dev.get(d_map, (capacity,), np.int32)[:k]
        if label is not None:
            out["label"] = dev.get(d_lab, (capacity,), np.int32)[:k]
    return out


def test_static_scratch_variant_equals_allocating_entry_point():
    case_a, case_b = make_case(600, 3, 42), make_case(400, 2, 7)
    W, H = case_a.width, case_a.height
    cam = (case_a.cx, case_a.cy, case_a.fx, case_a.fy)
    # (depth images, label mask or None, stride): different row counts and totals, so stale scratch would show
    two = np.stack([case_a.scene.source_depth.reshape(H, W), case_b.scene.source_depth.reshape(H, W)]).astype(np.int32)
    calls = [
        (case_a.scene.source_depth.reshape(1, H, W), case_a.scene.source_label.reshape(1, H, W), 4),
        (two, None, 8),
        (case_b.scene.source_depth.reshape(1, H, W), None, 1),
        (case_a.scene.source_depth.reshape(1, H, W), case_a.scene.source_label.reshape(1, H, W), 4),
    ]
    with Renderer(0) as r:
        dev = _Dev(r._lib, r._h)
        try:
            for depth, label, stride in calls:
                n_img = depth.shape[0]
                want = r.depth2cloud_global(depth.ravel(), None, W, H, n_img, *cam, 100.0, stride,
                                            None if label is None else label.ravel(), None, None)
                n_want = want["rendered_cloud_point_num"]
                assert n_want > 0
                got = _static_call(r, dev, depth, label, W, H, n_img, cam, stride, n_want + 7)
                assert got["rc"] == 0, r._lib.perch_last_error()
                assert got["n"] == n_want
                # bit-exact: same kernels, same order (the tolerance the depth2cloud parity tests use)
                assert np.array_equal(got["aos"].view(np.uint32), want["result_cloud_eigen"].view(np.uint32))
                assert np.array_equal(got["soa"].view(np.uint32), want["result_cloud"].view(np.uint32))
                assert np.array_equal(got["map"], want["cloud_pose_map"])
                if label is not None:
                    assert np.array_equal(got["label"], want["result_observed_cloud_label"])
                dev.free()
        finally:
            dev.free()


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    with Renderer(0) as r:
        lib, h = r._lib, r._h
        dev = _Dev(lib, h)
        try:
            d = dev.put(np.zeros(16, np.int32))
            n = C.c_int(0)
            # 4 x 4 images at stride 1 = 4 rows each: one image more than the array holds; rejected before any launch
            n_img = MAX_ROWS // 4 + 1
            rc = lib.perch_depth2cloud_static(h, d, None, 4, 4, n_img, 1.0, 1.0, 1.0, 1.0, 100.0, 1, 0, None, None, None, None,
                                              C.byref(n))
            assert rc == _capi.PERCH_ERR_CAPACITY
            assert b"static scratch" in lib.perch_last_error()
            # too small a capacity: count reported, error returned, nothing written past the buffers
            img = np.full((1, 4, 4), 50, np.int32)
            d2, d_aos = dev.put(img), dev.alloc(3 * 12)
            rc = lib.perch_depth2cloud_static(h, d2, None, 4, 4, 1, 1.0, 1.0, 1.0, 1.0, 100.0, 1, 3, d_aos, None, None, None,
                                              C.byref(n))
            assert rc == _capi.PERCH_ERR_CAPACITY and n.value == 16
        finally:
            dev.free()
