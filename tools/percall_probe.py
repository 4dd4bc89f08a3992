import sys, time
from pathlib import Path
ROOT = Path("/root/repo"); sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import bench
from perception_b200 import synth
r, wl = bench.build_workload("c2", 0, 0, None)
s = wl["scene"]
kw = dict(stride=wl["stride"], cost_type=wl["cost_type"], calculate_observed_cost=True, sensor_resolution=wl["res"])
P = 700
def T(f, n=20):
    f(); r.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): f()
    r.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("set_camera", T(lambda: r.set_camera(bench.W, bench.H, wl["proj"], synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)))
print("set_models", T(lambda: r.set_models(s.tris, s.tris_model_count)))
print("set_scene", T(lambda: r.set_scene(s.source_depth, s.source_color, s.source_label)))
print("set_observed", T(lambda: r.set_observed(wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"], wl["obs_label"])))
print("cost 700 (grid resident)", T(lambda: r.render_cuda_multi_unified("COST", wl["poses"][:P], wl["pmm"][:P], wl["plabel"][:P], wl["tot"][:P], **kw)))
def both():
    r.set_observed(wl["obs"]["result_cloud_eigen"], wl["obs"]["result_cloud_color"], wl["obs_label"])
    r.render_cuda_multi_unified("COST", wl["poses"][:P], wl["pmm"][:P], wl["plabel"][:P], wl["tot"][:P], **kw)
print("set_observed + cost 700 (grid rebuilt)", T(both))
def allc():
    r.set_camera(bench.W, bench.H, wl["proj"], synth.CX, synth.CY, synth.FX, synth.FY, synth.DEPTH_FACTOR)
    r.set_models(s.tris, s.tris_model_count)
    r.set_scene(s.source_depth, s.source_color, s.source_label)
    both()
print("all setters + cost 700", T(allc))
