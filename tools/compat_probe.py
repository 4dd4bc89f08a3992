"""The drop-in C++ backend (csrc/compat/cuda_renderer_compat.cpp, built against the reference's unchanged renderer.h into
oracle/_ref/libcuda_renderer_compat.so) timed the way the reference calls it: render_cuda_multi_unified("COST", ...) with
meshes, scene and observed cloud passed on EVERY call (std::vector / cv::Mat arguments built by the test trampolines),
in batches of the reference's gpu_batch_size 700 and in one call.  Wall clock per call, host side included.
Usage (GPU box): python tools/compat_probe.py"""
import ctypes as C
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
from perception_b200 import synth  # noqa: E402
from perception_b200.renderer import TRIANGLE_DTYPE  # noqa: E402

lib = C.CDLL(str(ROOT / "oracle/_ref/libcuda_renderer_compat.so"))
lib.compat_last_error.restype = C.c_char_p


def p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


r, wl = bench.build_workload("c2", 0, 0, None)
r.close()
s = wl["scene"]
tris = np.ascontiguousarray(s.tris, dtype=TRIANGLE_DTYPE)
cnt = np.ascontiguousarray(s.tris_model_count, np.int32)
rgb_in = np.ascontiguousarray(np.stack([s.source_color[k] for k in range(3)]).astype(np.uint8))
depth = np.ascontiguousarray(s.source_depth, np.int32)
label = np.ascontiguousarray(s.source_label.astype(np.uint8))
proj = np.ascontiguousarray(wl["proj"], np.float32)
obs_xyz = np.ascontiguousarray(wl["obs"]["result_cloud_eigen"], np.float32)
obs_rgb = np.ascontiguousarray(wl["obs"]["result_cloud_color"], np.uint8)
obs_lab = np.ascontiguousarray(wl["obs_label"], np.int32)


def call(lo, hi):
    P = hi - lo
    costs = np.full(3 * P, np.nan, np.float32)
    n = C.c_int(0)
    rc = lib.compat_render(b"COST", p(tris), len(tris), p(np.ascontiguousarray(wl["poses"][lo:hi], np.float32)), P,
                           p(np.ascontiguousarray(wl["pmm"][lo:hi], np.int32)), p(cnt), len(cnt), bench.W, bench.H, p(proj), p(depth),
                           p(rgb_in), p(label), p(np.ascontiguousarray(wl["plabel"][lo:hi], np.int32)), wl["stride"], 100,
                           C.c_float(synth.CX), C.c_float(synth.CY), C.c_float(synth.FX), C.c_float(synth.FY), p(obs_xyz), p(obs_rgb),
                           len(obs_xyz), p(np.ascontiguousarray(wl["tot"][lo:hi], np.float32)), p(obs_lab), wl["cost_type"], 1,
                           C.c_float(wl["res"]), C.c_float(20.0), C.c_float(1.0), None, None, 0, None, C.byref(n), p(costs))
    assert rc == 0, lib.compat_last_error()
    return costs


P = wl["P"]
for batch in (700, P):
    call(0, min(batch, P))  # warm-up: context, first mesh filing
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        for lo in range(0, P, batch):
            call(lo, min(lo + batch, P))
    dt = (time.perf_counter() - t0) / reps
    print(f"batch {batch}: {P / dt:,.0f} poses/s, {dt * 1e3 / ((P + batch - 1) // batch):.2f} ms per call, {dt * 1e3:.1f} ms per {P} poses")
